"""lr_ray_intersect_noalloc: the ray query with its scratch in the library's file-scope device array (no cudaMalloc).

No counterpart in the reference; what it must return is what
the existing entry point lr_ray_intersect returns (Scene::ray_intersect, src/render/scene.cpp:204-226), bit for bit, which is
the tolerance of every ray-cast test here (tests/test_gpu_parity.py).  Collected last on purpose (file name).
The test does not depend on state left by other tests -- it builds its own contexts."""
import numpy as np
import pytest

import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes

pytestmark = pytest.mark.gpu


def _rays(rng, n, lo, hi):
    o = rng.uniform(lo, hi, size=(n, 3)); t = rng.uniform(lo, hi, size=(n, 3))
    d = t - o; d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.zeros((n, 8), dtype=np.float32); rays[:, :3] = o; rays[:, 3:6] = d; rays[:, 6] = 3.0e38
    rays[: n // 4, 6] = rng.uniform(0.1, np.linalg.norm(hi - lo), size=n // 4)     # finite maxt
    rays[n // 4: n // 4 + 8, 3:6] = [1, 0, 0]                                       # axis-parallel rays (NaN slabs)
    return rays


def _same(a, b):
    return np.array_equal(a.view(np.uint32), b.view(np.uint32))


def test_noalloc_ray_intersect_matches_existing_entry_point_twice():
    cap = int(mi.Context(0).lib.lr_noalloc_max_rays())
    assert cap == 8192
    rng = np.random.default_rng(2026)
    liver = mi.Context(0).upload(mi.load_dict(scenes.liver_single_mesh(res=16, spp=1, grid_res=8)))
    cornell = mi.Context(0).upload(mi.load_dict(scenes.cornell_box_path(res=16, spp=1)))
    lo_l, hi_l = np.array([-65, -40, -65.]), np.array([-20, 5, -20.])
    lo_c, hi_c = np.array([-1.2] * 3), np.array([1.2] * 3)

    # two calls in a row on one context, different sizes (the second leaves stale rays / hits / stack entries of the first
    # in the shared array beyond its own n), then a call on ANOTHER context (same device, same array), then the first again
    r1 = _rays(rng, cap, lo_l, hi_l); r2 = _rays(rng, 1000, lo_l, hi_l); r3 = _rays(rng, 4097, lo_c, hi_c)
    ref1, ref2, ref3 = liver.ray_intersect(r1), liver.ray_intersect(r2), cornell.ray_intersect(r3)
    assert np.isfinite(ref1[:, 0]).sum() > cap // 10 and np.isfinite(ref3[:, 0]).sum() > 1000
    got1 = liver.ray_intersect_noalloc(r1)
    got2 = liver.ray_intersect_noalloc(r2)
    got3 = cornell.ray_intersect_noalloc(r3)
    got1b = liver.ray_intersect_noalloc(r1)
    assert _same(got1, ref1), f"first call: {(got1.view(np.uint32) != ref1.view(np.uint32)).any(axis=1).sum()} rays differ"
    assert _same(got2, ref2), f"second call: {(got2.view(np.uint32) != ref2.view(np.uint32)).any(axis=1).sum()} rays differ"
    assert _same(got3, ref3) and _same(got1b, ref1)

    # device buffers: nothing staged, the kernel is enqueued on the context's stream
    import torch
    d_r = torch.from_numpy(r2).cuda(); d_h = torch.full((r2.shape[0], 4), -1.0, device="cuda")
    torch.cuda.synchronize()
    liver.ray_intersect_noalloc(d_r, d_h); liver.ray_intersect_noalloc(d_r, d_h)
    torch.cuda.synchronize()
    assert _same(d_h.cpu().numpy(), ref2)

    # empty batch is a no-op; one ray more than the array holds is refused with a clear message, and nothing was launched
    assert liver.ray_intersect_noalloc(np.zeros((0, 8), np.float32)).shape == (0, 4)
    with pytest.raises(Exception, match="exceed the static scratch"):
        liver.ray_intersect_noalloc(_rays(rng, cap + 1, lo_l, hi_l))
    assert _same(liver.ray_intersect_noalloc(r2), ref2)
