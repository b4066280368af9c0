"""CPU tests (no GPU): the oracle against the reference's own end-to-end fixtures.
 - participating-media golden EXRs (src/render/tests/test_renders.py:181-270) with the reference's Z-test criterion
 - the Cornell-box crop pixel (src/integrators/tests/test_integrators.py:29-52)
 - film / reconstruction-filter closed forms (src/render/tests/test_imageblock.py, src/films/tests/test_hdrfilm.py)
Sized so the whole CPU suite stays within a few minutes on 8 cores."""
import os
import numpy as np
import pytest

import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes
from tests import orc, golden_scenes as G


@pytest.mark.parametrize("name", G.GOLDEN_SCENES)
def test_oracle_passes_reference_ztest(name):
    """Same acceptance as the reference's test_render: per-pixel Z-test (XYZ, variance floor 1e-4,
    Sidak alpha) with >= 99.75 % of pixels accepted.  The reference renders 2e6/(48*48) = 868 spp."""
    spp = int(2e6) // (48 * 48)
    sc = mi.load_dict(G.scene(name, spp=spp))
    img = orc.OracleScene(sc).render(seed=0, spp=spp)
    frac, ok = G.reference_accepts(img, name, spp)
    assert ok, f"{name}: only {frac:.4f} of the entries accepted"


def test_oracle_prb_primal_passes_reference_ztest():
    """INTEGRATOR_MAPPING (test_renders.py:45-49): *_volpath scenes are also rendered with prbvolpath."""
    spp = int(2e6) // (48 * 48); name = "test_hetero_sigmat"
    sc = mi.load_dict(G.scene(name, integrator="prbvolpath", spp=spp))
    img = orc.OracleScene(sc).render(seed=0, spp=spp)
    frac, ok = G.reference_accepts(img, name, spp)
    assert ok, f"{name} (prbvolpath): only {frac:.4f} accepted"


def test_cornell_directly_visible_pixel():
    """test_integrators.py:29-52: pixel (124, 36) of mi.cornell_box() looks at the light; path max_depth=1
    => exactly the emitter radiance; hide_emitters => 0.  (The crop is restated as lanes of that pixel.)"""
    d = mi.cornell_box(); d["integrator"] = {"type": "path", "max_depth": 1, "hide_emitters": False}
    sc = mi.load_dict(d); spp = 16; pix = 36 * 256 + 124
    lanes = orc.OracleScene(sc).sample_lanes(0, spp, pix * spp, (pix + 1) * spp)
    assert np.allclose(lanes[:, :3], [18.387, 13.9873, 6.75357], rtol=1e-5)
    d["integrator"]["hide_emitters"] = True
    lanes = orc.OracleScene(mi.load_dict(d)).sample_lanes(0, spp, pix * spp, (pix + 1) * spp)
    assert np.all(lanes[:, :3] == 0)


def test_box_film_is_mean_of_lanes():
    """hdrfilm + box filter (imageblock.cpp:206-235, hdrfilm.cpp:402): pixel = mean of its lanes' radiance."""
    sc = mi.load_dict(scenes.liver_single_mesh(res=8, spp=4, grid_res=16)); o = orc.OracleScene(sc)
    img = o.render(seed=3, spp=4); lanes = o.sample_lanes(3, 4, 0, 8 * 8 * 4)
    ref = lanes[:, :3].reshape(8, 8, 4, 3).mean(axis=2)
    assert np.allclose(img[..., :3], ref, rtol=1e-5, atol=1e-7)
    assert np.allclose(img[..., 3], lanes[:, 3].reshape(8, 8, 4).mean(axis=2), atol=1e-6)     # alpha = mean(valid)


def test_gaussian_film_weights_normalise():
    """A constant-radiance image stays constant under the gaussian splat + develop (weights cancel)."""
    d = mi.cornell_box(); d["integrator"] = {"type": "path", "max_depth": 1}
    d["sensor"]["film"]["width"] = d["sensor"]["film"]["height"] = 16
    d["env"] = {"type": "constant", "radiance": {"type": "rgb", "value": [0.25, 0.5, 1.0]}}
    for k in ("floor", "ceiling", "back", "green-wall", "red-wall", "small-box", "large-box", "light"):
        d.pop(k)
    d["far"] = {"type": "rectangle", "to_world": mi.T().translate([0, 0, 50]), "bsdf": {"type": "diffuse"}}   # behind the camera
    img = orc.OracleScene(mi.load_dict(d)).render(seed=0, spp=8)
    assert np.allclose(img, np.broadcast_to(np.float32([0.25, 0.5, 1.0]), img.shape), rtol=2e-5)


def test_lane_split_equals_whole_render():
    """Streams are keyed by the global lane index: rendering two halves into one film == one render."""
    sc = mi.load_dict(scenes.cornell_box_path(res=16, spp=4)); n = 16 * 16 * 4
    o = orc.OracleScene(sc); whole = o.render(seed=5, spp=4)
    o.film_clear(); o.render_lanes(5, 4, 0, n // 2); o.render_lanes(5, 4, n // 2, n)
    assert np.allclose(o.film_develop(), whole, rtol=1e-5, atol=1e-7)
