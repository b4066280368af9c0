"""One frame split over devices by sample index (lr_set_lane_map, multi.sample_share): what N ranks render is the single-device
render.  The reference has no multi-GPU code to match; the property tested is the north star's "partitioned ... by image tile /
sample index with no communication in the forward pass": PCG32 streams are keyed by the GLOBAL lane
(src/render/integrator.cpp:321-338, sampler.cpp:129-148), so per-lane radiance is bit-identical whichever rank traces it, film
WEIGHTS add up exactly and RGB adds up to the order of fp32 atomics.  Runs on one device (two contexts stand for two ranks) and,
when the box has two, with one context per device."""
import numpy as np
import pytest

import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes, multi

pytestmark = pytest.mark.gpu


def _n_devices():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("tracking", ["dda", "reference"])
def test_mapped_lanes_are_the_same_lanes(tracking):
    """sample_lanes under a lane map returns, row for row, the rows of the unmapped call (wavefront volpath and per-lane path)."""
    for d, res, spp in ((scenes.liver_single_mesh(res=24, spp=8, grid_res=16, tracking=tracking), 24, 8), (scenes.cornell_box_path(res=16, spp=6), 16, 6)):
        ctx = mi.Context(0).upload(mi.load_dict(d)); n = res * res * spp
        full = ctx.sample_lanes(7, spp, 0, n)
        for world in (2, 3):
            for rank in range(world):
                share = multi.sample_share(spp, rank, world); ctx.set_lane_map(*share)
                n_loc = n // spp * share[2]
                part = ctx.sample_lanes(7, spp, 0, n_loc)
                sub = ctx.sample_lanes(7, spp, 5, 77)                       # a sub-range of local indices
                ctx.set_lane_map()
                g = multi.lane_global(share, np.arange(n_loc))
                assert np.array_equal(part.view(np.uint32), full[g].view(np.uint32)), (world, rank)
                assert np.array_equal(sub.view(np.uint32), full[g[5:77]].view(np.uint32))
        with pytest.raises(RuntimeError, match="lr_set_lane_map"):
            ctx.set_lane_map(spp, spp - 1, 2)
        ctx.close()


def _split_film(devices, d, seed, spp):
    world = len(devices); films = []
    for rank, dev in enumerate(devices):
        ctx = mi.Context(dev).upload(mi.load_dict(d))
        multi.render_distributed(ctx, seed, spp, rank, world, mode="samples")      # no reduction inside: films are summed below
        films.append(ctx.film_read().astype(np.float64)); ctx.close()
    return sum(films)


def _check_split(devices):
    res, spp = 48, 32
    d = scenes.liver_single_mesh(res=res, spp=spp, grid_res=32, tracking="dda")
    ctx = mi.Context(devices[0]).upload(mi.load_dict(d)); ctx.film_clear(); ctx.render_lanes(3, spp, 0, 0); one = ctx.film_read().astype(np.float64); ctx.close()
    both = _split_film(devices, d, 3, spp)
    assert np.array_equal(both[..., -1], one[..., -1]) and one[..., -1].min() == spp          # weights: exact
    assert np.allclose(both[..., :3], one[..., :3], rtol=2e-5, atol=1e-5)                      # radiance: the order of fp32 atomic adds
    # the adjoint: gradient shares add up to the whole
    di = scenes.sphere_liver_inverse(res=32, spp=16, grid_res=16, tracking="dda"); g_in = np.ones((32, 32, 3), np.float32)
    c0 = mi.Context(devices[0]).upload(mi.load_dict(di)); c0.grid_grad_zero(0); c0.render_backward(g_in, seed=5, spp=16); whole = c0.grid_grad_read(0).astype(np.float64); c0.close()
    parts = []
    for rank, dev in enumerate(devices):
        c = mi.Context(dev).upload(mi.load_dict(di)); parts.append(multi.backward_distributed(c, g_in, 5, 16, rank, len(devices)).astype(np.float64)); c.close()
    tot = sum(parts); scale = np.abs(whole).max()
    assert scale > 0 and np.abs(tot - whole).max() <= 2e-4 * scale, (np.abs(tot - whole).max(), scale)
    assert all(np.abs(p).max() > 0 for p in parts)


def test_sample_split_film_and_gradient_one_device():
    _check_split([0, 0]); _check_split([0, 0, 0])


def test_sample_split_film_and_gradient_two_devices():
    if _n_devices() < 2:
        pytest.skip("needs two CUDA devices")
    _check_split([0, 1])


def test_two_contexts_with_different_trees_share_the_trace_kernels():
    """The dynamic shared-memory limit of the TRACE kernels is a per-function attribute: uploading a scene with a SMALL tree in a second
    context must not stop the first context (larger tree, more shared memory) from rendering."""
    big = mi.Context(0).upload(mi.load_dict(scenes.sphere_liver_inverse(res=32, spp=4, grid_res=16, tracking="dda")))      # 20 480 triangles
    a = big.render(seed=1, spp=4)
    small = mi.Context(0).upload(mi.load_dict(scenes.cornell_box_fog(res=32, spp=4)))                                      # a few dozen
    small.render(seed=1, spp=4)
    b = big.render(seed=1, spp=4)
    assert np.allclose(a, b, rtol=1e-5, atol=1e-7)                     # the same lanes; film atomics in another order
