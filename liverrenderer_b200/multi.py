"""Host-side multi-GPU helpers: one process per GPU, scene replicated, lanes independent.

Forward: no communication while rendering; one reduction of the film accumulator at the end.
Inverse: one all-reduce of the grid gradient per step.  Works over any `torch.distributed` backend
(`nccl` with CUDA tensors on the B200 box, `gloo` with CPU tensors in the CPU tests)."""
from __future__ import annotations

import numpy as np


def lane_count(width: int, height: int, spp: int) -> tuple[int, int, int]:
    """(lanes per pass, spp per pass, passes): the reference's pass-splitting rule when W*H*spp exceeds 2^32-1
    (src/render/integrator.cpp:275-293)."""
    wf = width * height * spp; limit = 0xffffffff
    spp_per_pass, n_passes = spp, 1
    if wf > limit:
        spp_per_pass = spp // ((wf + limit - 1) // limit); n_passes = spp // spp_per_pass
    return width * height * spp_per_pass, spp_per_pass, n_passes


def lane_range(n_lanes: int, rank: int, world: int, align: int = 32) -> tuple[int, int]:
    """Contiguous share of the global lane range [0, n_lanes) for `rank`; boundaries aligned to warps.
    Streams are keyed by the GLOBAL lane index, so the union of all ranks' ranges reproduces the single-GPU render."""
    per = -(-n_lanes // world); per = -(-per // align) * align
    b = min(rank * per, n_lanes); e = min(b + per, n_lanes)
    return b, e


def interleaved_blocks(n_lanes: int, rank: int, world: int, block: int = 1 << 16):
    """Round-robin blocks of lanes (load balance: expensive pixels are spread over ranks)."""
    k = rank
    while k * block < n_lanes:
        yield k * block, min((k + 1) * block, n_lanes)
        k += world


def allreduce_sum_(t, group=None):
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, group=group)
    return t


def develop(film):
    """HDRFilm::develop (src/films/hdrfilm.cpp:402): out = value / (W == 0 ? 1 : W); works on numpy or torch [H,W,C]."""
    w = film[..., -1:]
    return film[..., :-1] / (w + (w == 0))


def render_distributed(ctx, seed: int, spp: int, rank: int, world: int, mode: str = "lanes", film_buf=None):
    """Render this rank's share into ctx's film, reduce the film over ranks, return the developed image buffer
    (`film_buf`: torch tensor [H,W,C] on the device the backend uses for communication)."""
    W, H = ctx.scene.film_size
    ctx.film_clear()
    if mode == "lanes":
        n, _, _ = lane_count(W, H, spp)
        for b, e in interleaved_blocks(n, rank, world):
            ctx.render_lanes(seed, spp, b, e)
    else:                                   # "seeds": every rank renders everything with its own seed
        ctx.render_lanes(seed + rank, spp, 0, 0)
    if film_buf is not None:
        ctx.film_read(film_buf); allreduce_sum_(film_buf); ctx.film_write(film_buf)
    return ctx.film_develop()


def backward_distributed(ctx, grad_in, seed: int, spp: int, rank: int, world: int, grid: int = 0, grad_buf=None):
    """Adjoint over this rank's lanes + all-reduce of the grid gradient."""
    W, H = ctx.scene.film_size
    n, _, _ = lane_count(W, H, spp); b, e = lane_range(n, rank, world)
    ctx.grid_grad_zero(grid)
    if e > b:
        ctx.render_backward(grad_in, seed=seed, spp=spp, begin=b, end=e)
    if grad_buf is not None:
        ctx.grid_grad_read(grid, grad_buf); allreduce_sum_(grad_buf)
        return grad_buf
    return ctx.grid_grad_read(grid)
