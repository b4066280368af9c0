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


def test_cornell_crop_window_as_the_reference_tests_it():
    """test_integrators.py:29-52 as written there: an actual crop (crop_offset 124, 36; crop 1 x 1) of mi.cornell_box(), path max_depth = 1:
    the rendered 1 x 1 image is the emitter radiance; hide_emitters => 0.  And the crop is the SAME camera: a crop window renders what
    the full frame shows there (identical sub-pixel jitter => identical rays: crop pixel (x, y) with lane-local jitter j looks through
    full-frame position (x + ox + j) -- checked on a window that straddles the light's edge)."""
    d = mi.cornell_box(); d["integrator"] = {"type": "path", "max_depth": 1, "hide_emitters": False}
    d["sensor"]["film"].update({"crop_offset_x": 124, "crop_offset_y": 36, "crop_width": 1, "crop_height": 1})
    sc = mi.load_dict(d); assert sc.film_size == (1, 1)
    img = orc.OracleScene(sc).render(seed=0, spp=64)
    assert img.shape[:2] == (1, 1) and np.allclose(img[0, 0, :3], [18.387, 13.9873, 6.75357], rtol=1e-5)
    d["integrator"]["hide_emitters"] = True
    assert np.all(orc.OracleScene(mi.load_dict(d)).render(seed=0, spp=64)[..., :3] == 0)
    # window vs full frame: the directly visible emitter (max_depth 1) is a deterministic function of the ray
    full = mi.cornell_box(); full["integrator"] = {"type": "path", "max_depth": 1}; full["sensor"]["film"]["rfilter"] = {"type": "box"}
    win = mi.cornell_box(); win["integrator"] = dict(full["integrator"]); win["sensor"]["film"]["rfilter"] = {"type": "box"}
    ox, oy, w, h = 96, 24, 48, 32
    win["sensor"]["film"].update({"crop_offset_x": ox, "crop_offset_y": oy, "crop_width": w, "crop_height": h})
    spp = 256
    a = orc.OracleScene(mi.load_dict(full)).render(seed=1, spp=spp)[oy:oy + h, ox:ox + w, :3]
    b = orc.OracleScene(mi.load_dict(win)).render(seed=2, spp=spp)[..., :3]
    assert b.shape == (h, w, 3)
    lit_a, lit_b = a[..., 0] > 0, b[..., 0] > 0
    inside = (a[..., 0] == a.max()) & (b[..., 0] == b.max()); outside = ~lit_a & ~lit_b
    assert inside.sum() > 100 and outside.sum() > 100 and (inside | outside).mean() > 0.9      # same silhouette but for the edge pixels
    edge = ~(inside | outside); assert np.abs(a[edge] - b[edge]).max() < 0.2 * a.max()            # coverage of an edge pixel: MC at 256 spp
    with pytest.raises(RuntimeError, match="crop"):
        bad = mi.cornell_box(); bad["sensor"]["film"].update({"crop_offset_x": 250, "crop_width": 10}); mi.load_dict(bad)


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


def _albedo_grid_scene(res, spp, tracking="reference"):
    """A scattering cube whose ALBEDO is an rgb gridvolume (heterogeneous.cpp:177-186 exposes `albedo` and `sigma_t` as differentiable)."""
    rng = np.random.default_rng(4)
    alb = (0.3 + 0.6 * rng.random((4, 4, 4, 3))).astype(np.float32); sig = (0.4 + 0.6 * rng.random((4, 4, 4, 1))).astype(np.float32)
    T = mi.T
    return {"type": "scene", "integrator": {"type": "prbvolpath", "max_depth": 6, "rr_depth": 64, "tracking": tracking},
            "sensor": {"type": "perspective", "fov": 30.0, "near_clip": 0.01, "far_clip": 100.0,
                       "to_world": T().look_at(origin=[0.3, 0.4, -4.0], target=[0, 0, 0], up=[0, 1, 0]),
                       "sampler": {"type": "independent", "sample_count": spp},
                       "film": {"type": "hdrfilm", "width": res, "height": res, "rfilter": {"type": "box"}, "pixel_format": "rgb"}},
            "cube": {"type": "cube", "to_world": T().scale(0.5), "bsdf": {"type": "null"},
                     "interior": {"type": "heterogeneous", "scale": 6.0,
                                  "sigma_t": {"type": "gridvolume", "grid": sig, "max_value": 1.5, "to_world": T().translate([-0.5, -0.5, -0.5])},
                                  "albedo": {"type": "gridvolume", "grid": alb, "to_world": T().translate([-0.5, -0.5, -0.5])}}},
            "env": {"type": "constant", "radiance": {"type": "rgb", "value": [1, 1, 1]}}}


def test_albedo_grid_gradient_against_finite_differences():
    """d loss / d albedo-grid (d weight_c / d a_c = s * sigma_t / P_s; prbvolpath.py:202-204 through sigma_s = sigma_t * albedo).
    The albedo never enters a sampling decision (roulette off), so for fixed PCG32 streams the image is a POLYNOMIAL of the albedo
    values: central differences with the same seed carry no Monte-Carlo noise, only the O(eps^2) curvature term."""
    res, spp = 12, 64
    sc = mi.load_dict(_albedo_grid_scene(res, spp)); o = orc.OracleScene(sc)
    ga = sc.param_keys["cube.interior_medium.albedo.data"]; assert sc.grids[ga].differentiable == 1 and sc.grids[ga].channels == 3
    rng = np.random.default_rng(9); w = rng.random((res, res, 3)).astype(np.float32)
    o.grid_grad_zero(ga); o.render_backward(w, seed=21, spp=spp); grad = o.grid_grad_read(ga).reshape(4, 4, 4, 3).astype(np.float64)
    assert np.abs(grad).max() > 0
    base = np.array(sc.grid_arrays[ga], dtype=np.float32)
    def loss(a):
        o.grid_update(ga, a); return float((o.render(seed=21, spp=spp).astype(np.float64) * w).sum())
    eps = 2e-2
    for c in range(3):                                               # a uniform shift of one channel
        d = np.zeros_like(base); d[..., c] = eps
        fd = (loss(base + d) - loss(base - d)) / (2 * eps)
        assert abs(fd - grad[..., c].sum()) <= 2e-3 * abs(fd) + 1e-6, (c, fd, grad[..., c].sum())
    d = np.zeros_like(base); d[1, 2, 1, 0] = eps                     # a single voxel
    fd = (loss(base + d) - loss(base - d)) / (2 * eps)
    assert abs(fd - grad[1, 2, 1, 0]) <= 5e-3 * abs(fd) + 1e-6, (fd, grad[1, 2, 1, 0])
    o.grid_update(ga, base)
