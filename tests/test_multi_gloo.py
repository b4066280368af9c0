"""CPU tests of the multi-GPU host logic with world_size 2 over gloo (no GPU): lane partitions cover the wavefront
exactly once, film reduction + develop equals the single-process film, gradient all-reduce sums.  The per-rank
"renderer" here is the oracle (test infrastructure) standing in for a GPU context."""
import os
import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from liverrenderer_b200 import multi


def test_lane_ranges_partition_exactly():
    for n in (1, 31, 32, 1000, 4194304, 1080 * 1080 * 1024):
        for world in (1, 2, 3, 4, 8):
            covered = 0; last = 0
            for r in range(world):
                b, e = multi.lane_range(n, r, world)
                assert b == last or b == e; assert b % 32 == 0 or b == n
                covered += e - b; last = max(last, e)
            assert covered == n and last == n
            blocks = sorted(x for r in range(world) for x in multi.interleaved_blocks(n, r, world, block=4096))
            assert blocks[0][0] == 0 and blocks[-1][1] == n and all(a[1] == b[0] for a, b in zip(blocks, blocks[1:]))


def test_pass_splitting_rule():
    """src/render/integrator.cpp:275-293: C4 = 1080x1080x4096 spp > 2^32 lanes => 2 passes of 2048 spp."""
    assert multi.lane_count(1080, 1080, 1024) == (1080 * 1080 * 1024, 1024, 1)
    n, spp_pp, passes = multi.lane_count(1080, 1080, 4096)
    assert (spp_pp, passes) == (2048, 2) and n == 1080 * 1080 * 2048 and n <= 0xffffffff


def test_sample_shares_partition_exactly():
    """multi.sample_share: the ranks' sample intervals tile [0, spp) and the local <-> global lane maps are inverse bijections."""
    for spp in (1, 3, 4, 64, 1024, 2048):
        for world in (1, 2, 3, 8):
            shares = [multi.sample_share(spp, r, world) for r in range(world)]
            cover = sorted((s[1], s[1] + s[2]) for s in shares if s is not None)
            assert cover[0][0] == 0 and cover[-1][1] == spp and all(a[1] == b[0] for a, b in zip(cover, cover[1:]))
            assert all(s is None or s[0] == spp for s in shares)
            n_px = 5; seen = []
            for s in shares:
                if s is None:
                    continue
                loc = np.arange(n_px * s[2]); g = multi.lane_global(s, loc)
                assert np.array_equal(multi.lane_local(s, g), loc) and np.all((g % spp >= s[1]) & (g % spp < s[1] + s[2]))
                seen.append(g)
            assert np.array_equal(np.sort(np.concatenate(seen)), np.arange(n_px * spp))


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import liverrenderer_b200 as mi
    from liverrenderer_b200 import scenes
    from tests import orc
    sc = mi.load_dict(scenes.liver_single_mesh(res=12, spp=4, grid_res=8)); o = orc.OracleScene(sc, threads=1)
    n, _, _ = multi.lane_count(12, 12, 4)
    o.film_clear()
    for b, e in multi.interleaved_blocks(n, rank, world, block=64):
        o.render_lanes(3, 4, b, e)
    film = torch.from_numpy(o.film_read()); multi.allreduce_sum_(film)
    img = multi.develop(film).numpy()
    # inverse: each rank's half of the lanes, gradient all-reduced
    sc2 = mi.load_dict(scenes.sphere_liver_inverse(res=8, spp=4, grid_res=4)); o2 = orc.OracleScene(sc2, threads=1)
    n2, _, _ = multi.lane_count(8, 8, 4); b, e = multi.lane_range(n2, rank, world)
    g_in = np.ones((8, 8, 3), dtype=np.float32); o2.grid_grad_zero(0); o2.render_backward(g_in, seed=5, spp=4, begin=b, end=e)
    grad = torch.from_numpy(o2.grid_grad_read(0)); multi.allreduce_sum_(grad)
    # the sample-index split of one frame (multi.sample_share): every rank takes spp / world samples of every pixel
    share = multi.sample_share(4, rank, world); o.film_clear()
    for px in range(12 * 12):
        o.render_lanes(3, 4, px * 4 + share[1], px * 4 + share[1] + share[2])
    film = torch.from_numpy(o.film_read()); multi.allreduce_sum_(film)
    img_s = multi.develop(film).numpy()
    if rank == 0:
        q.put((img, grad.numpy(), img_s))
    dist.barrier(); dist.destroy_process_group()


def test_world_size_2_film_and_gradient_reduce():
    import liverrenderer_b200 as mi
    from liverrenderer_b200 import scenes
    from tests import orc
    ctx = mp.get_context("spawn"); q = ctx.Queue(); port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    img, grad, img_s = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120); assert p.exitcode == 0
    sc = mi.load_dict(scenes.liver_single_mesh(res=12, spp=4, grid_res=8)); ref = orc.OracleScene(sc, threads=1).render(seed=3, spp=4)
    assert np.allclose(img, ref, rtol=1e-5, atol=1e-7) and np.allclose(img_s, ref, rtol=1e-5, atol=1e-7)
    sc2 = mi.load_dict(scenes.sphere_liver_inverse(res=8, spp=4, grid_res=4)); o2 = orc.OracleScene(sc2, threads=1)
    o2.grid_grad_zero(0); o2.render_backward(np.ones((8, 8, 3), dtype=np.float32), seed=5, spp=4)
    assert np.allclose(grad, o2.grid_grad_read(0), rtol=1e-4, atol=1e-9)
