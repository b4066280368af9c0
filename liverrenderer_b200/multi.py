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


def sample_share(spp_per_pass: int, rank: int, world: int):
    """(period, offset, length) for `Context.set_lane_map`: rank `rank` of `world` renders samples [offset, offset + length) of
    every pixel (lanes of a pixel are consecutive: lane = pixel * spp_per_pass + sample, src/render/integrator.cpp:321-338).
    Every rank gets the same pixels, so the split is balanced whatever the image looks like.  None when the rank has no samples."""
    off = rank * spp_per_pass // world; length = (rank + 1) * spp_per_pass // world - off
    return (spp_per_pass, off, length) if length > 0 else None


def lane_global(share, i):
    """Local index -> global lane under a lane map (period, offset, length); mirrors `lane_global` in csrc/lr_kernels.cuh."""
    period, off, length = share
    return (i // length) * period + off + (i % length)


def lane_local(share, lane):
    period, off, length = share
    return (lane // period) * length + (lane % period - off)


class _DevicePtr:
    """Zero-copy view of a device allocation of the backend for torch (`torch.as_tensor(_DevicePtr(...), device=...)`)."""
    def __init__(self, ptr: int, shape, typestr="<f4"):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2, "strides": None}


def film_tensor(ctx, device=None):
    """The film accumulator [H, W, C] of `ctx` as a torch CUDA tensor aliasing `lr_film_device_ptr` (no copy): collectives run on it in place."""
    import torch
    W, H = ctx.scene.film_size
    return torch.as_tensor(_DevicePtr(ctx.film_device_ptr(), (H, W, ctx.film_channels())), device=device if device is not None else torch.device("cuda", ctx.device))


def grid_grad_tensor(ctx, grid: int = 0, device=None):
    """The gradient buffer of grid `grid` as a flat torch CUDA tensor aliasing `lr_grid_grad_device_ptr`."""
    import torch
    return torch.as_tensor(_DevicePtr(ctx.grid_grad_device_ptr(grid), (ctx.grid_size(grid),)), device=device if device is not None else torch.device("cuda", ctx.device))


def render_distributed(ctx, seed: int, spp: int, rank: int, world: int, mode: str = "samples", film_buf=None, out=None):
    """Render this rank's share into ctx's film, reduce the film over ranks, return the developed image.
    mode "samples": ONE frame, rank r takes spp / world samples of every pixel (`sample_share`); "lanes": ONE frame, round-robin
    blocks of lanes; "seeds": every rank renders the whole frame with its own seed (independent replicas of the image).
    `film_buf`: None = no reduction (single process); True = all-reduce in place on the film accumulator (CUDA + NCCL);
    a tensor [H, W, C] = staging buffer on the device the process group communicates on (CPU tensors for gloo)."""
    W, H = ctx.scene.film_size
    ctx.film_clear()
    if mode == "samples":
        _, spp_pass, _ = lane_count(W, H, spp)
        share = sample_share(spp_pass, rank, world)
        if share is not None:
            ctx.set_lane_map(*share)
            try:
                ctx.render_lanes(seed, spp, 0, 0)
            finally:
                ctx.set_lane_map()
    elif mode == "lanes":
        n, _, _ = lane_count(W, H, spp)
        for b, e in interleaved_blocks(n, rank, world):
            ctx.render_lanes(seed, spp, b, e)
    else:                                   # "seeds": every rank renders everything with its own seed
        ctx.render_lanes(seed + rank, spp, 0, 0)
    if film_buf is True:
        allreduce_sum_(film_tensor(ctx))
    elif film_buf is not None:
        ctx.film_read(film_buf); allreduce_sum_(film_buf); ctx.film_write(film_buf)
    return ctx.film_develop(out)


def backward_distributed(ctx, grad_in, seed: int, spp: int, rank: int, world: int, grid: int = 0, grad_buf=None, mode: str = "samples"):
    """Adjoint over this rank's lanes + all-reduce of the grid gradient.  `grad_buf`: None = no reduction; True = in place on the
    gradient buffer of the backend (CUDA + NCCL), returns that tensor; a tensor = staging buffer (CPU tensors for gloo)."""
    W, H = ctx.scene.film_size
    n, spp_pass, _ = lane_count(W, H, spp)
    ctx.grid_grad_zero(grid)
    if mode == "samples":
        share = sample_share(spp_pass, rank, world)
        if share is not None:
            ctx.set_lane_map(*share)
            try:
                ctx.render_backward(grad_in, seed=seed, spp=spp)
            finally:
                ctx.set_lane_map()
    else:
        b, e = lane_range(n, rank, world)
        if e > b:
            ctx.render_backward(grad_in, seed=seed, spp=spp, begin=b, end=e)
    if grad_buf is True:
        return allreduce_sum_(grid_grad_tensor(ctx, grid))
    if grad_buf is not None:
        ctx.grid_grad_read(grid, grad_buf); allreduce_sum_(grad_buf)
        return grad_buf
    return ctx.grid_grad_read(grid)
