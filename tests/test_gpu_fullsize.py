"""Full-size GPU checks at BASELINE.json's configurations, through size-independent properties (the oracle cannot
finish these sizes): every lane splats exactly once (the film's weight channel is a checksum of the lane hand-out and of
the wavefront batching / pool recycling), lane ranges add up to the whole render, pass splitting beyond 2^32 lanes
(src/render/integrator.cpp:275-293), and agreement of the converged image with a low-spp render of the same scene."""
import numpy as np
import pytest

import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes

pytestmark = pytest.mark.gpu


def test_c3_full_size_weight_checksum_and_lane_ranges():
    """C3 (Liver-SingleMesh, 256^3 heterogeneous gridvolume, volpath, 1080x1080, 1024 spp = 1 194 393 600 lanes, DDA tracking):
    box filter => each lane adds exactly 1.0 to one pixel's weight and alpha; 1024 is exact in fp32."""
    W = H = 1080; spp = 1024
    sc = mi.load_dict(scenes.liver_single_mesh(res=W, spp=spp, grid_res=256, tracking="dda"))
    ctx = mi.Context(0).upload(sc)
    ctx.film_clear(); ctx.render_lanes(3, spp, 0, 0)
    st = ctx.stats(); film = ctx.film_read()
    assert st["samples"] == W * H * spp
    assert film.shape == (H, W, 5) and np.isfinite(film).all()
    assert np.array_equal(film[..., 4], np.full((H, W), float(spp), np.float32))          # weight: every lane exactly once
    assert np.array_equal(film[..., 3], film[..., 4])                                       # environment emitter: every ray valid
    assert st["rays"] >= st["samples"] and st["collisions"] > 0
    whole = film[..., :3] / spp
    # the same lanes as three ranges cut at arbitrary (non pool-aligned) points: identical paths, fp32 atomics reordered
    n = W * H * spp; cuts = [0, n // 3 + 12345, 2 * (n // 3) + 7, n]
    ctx.film_clear()
    for a, b in zip(cuts[:-1], cuts[1:]):
        ctx.render_lanes(3, spp, a, b)
    film2 = ctx.film_read()
    assert np.array_equal(film2[..., 4], film[..., 4])
    parts = film2[..., :3] / spp
    assert np.allclose(parts, whole, rtol=2e-4, atol=1e-5)
    # converged image vs an independent 16-spp render: relative difference of the image means within 1 %
    ctx.film_clear(); ctx.render_lanes(99, 16, 0, 0); low = ctx.film_read()[..., :3] / 16
    assert abs(low.mean() - whole.mean()) / whole.mean() < 1e-2
    ctx.close()


def test_c4_full_size_two_passes():
    """C4 (Liver-MultiMesh, principled shell + roughdielectric inner boundary, two nested heterogeneous media, 1080x1080,
    4096 spp = 4 777 574 400 lanes > 2^32 - 1 => 2 passes of 2048 spp, each lane continuing its PCG32 stream)."""
    W = H = 1080; spp = 4096
    sc = mi.load_dict(scenes.liver_multi_mesh(res=W, spp=spp, tracking="dda"))
    ctx = mi.Context(0).upload(sc)
    ctx.film_clear(); ctx.render_lanes(1, spp, 0, 0)
    st = ctx.stats(); film = ctx.film_read()
    assert st["samples"] == W * H * spp
    assert np.isfinite(film).all()
    assert np.array_equal(film[..., -1], np.full((H, W), float(spp), np.float32))
    img = film[..., :3] / spp
    ctx.film_clear(); ctx.render_lanes(7, 32, 0, 0); low = ctx.film_read()[..., :3] / 32
    assert abs(low.mean() - img.mean()) / img.mean() < 1e-2
    # per-pixel: the 32-spp render scatters around the converged one; 8x8-binned means agree within 5 %
    b = lambda x: x.reshape(H // 8, 8, W // 8, 8, 3).mean(axis=(1, 3))
    hi, lo = b(img), b(low)
    rel = np.abs(hi - lo) / np.maximum(hi, 1e-2)
    assert np.quantile(rel, 0.99) < 0.12, float(np.quantile(rel, 0.99))
    ctx.close()


def test_c3_full_size_bumpmap_boundary():
    """C3 at full size with the fork's own boundary (bumpmap(bitmap) over a smooth dielectric):
    weight checksum, and the converged image against an independent low-spp render and against the same scene with the
    bump scale set to zero (a 0.005-scale bump map must not change the image mean by more than a few percent, but must
    change the image)."""
    W = H = 1080; spp = 256
    d = scenes.liver_single_mesh(res=W, spp=spp, grid_res=256, tracking="dda", surface="bumpmap")
    ctx = mi.Context(0).upload(mi.load_dict(d))
    ctx.film_clear(); ctx.render_lanes(5, spp, 0, 0); st = ctx.stats(); film = ctx.film_read()
    assert st["samples"] == W * H * spp and np.isfinite(film).all()
    assert np.array_equal(film[..., 4], np.full((H, W), float(spp), np.float32))
    img = film[..., :3] / spp
    ctx.film_clear(); ctx.render_lanes(11, 16, 0, 0); low = ctx.film_read()[..., :3] / 16
    assert abs(low.mean() - img.mean()) / img.mean() < 1e-2
    ctx.close()
    d["liver"]["bsdf"]["scale"] = 0.0
    c0 = mi.Context(0).upload(mi.load_dict(d)); c0.film_clear(); c0.render_lanes(5, spp, 0, 0); flat = c0.film_read()[..., :3] / spp; c0.close()
    assert abs(flat.mean() - img.mean()) / img.mean() < 5e-2
    b = lambda x: x.reshape(H // 8, 8, W // 8, 8, 3).mean(axis=(1, 3))
    assert np.quantile(np.abs(b(flat) - b(img)) / np.maximum(b(img), 1e-2), 0.999) > 1e-2      # the perturbed normals are visible
