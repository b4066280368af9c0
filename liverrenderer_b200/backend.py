"""Thin object wrapper over the C ABI (include/lr_api.h).

`Context` owns one `lr_ctx` (one CUDA device).  All compute goes through
liblr_b200.so; a missing library or missing GPU raises -- there is no CPU path.
"""
from __future__ import annotations

import ctypes as C
import numpy as np

from . import _capi as K
from .scene import Scene


def _ptr(a, n: int, what: str, convert: bool = False):
    """numpy array or torch tensor -> (void*, is_device, keep-alive).  The C side moves exactly `n` fp32 elements through the
    pointer, so the buffer must be float32, C-contiguous and hold exactly `n` elements.  `convert` (INPUT buffers only): anything
    else of the right element count is converted to a contiguous float32 copy; an OUTPUT buffer that does not qualify raises
    (TypeError for the dtype, ValueError for layout / size) -- writing through a reinterpreted or short buffer is never right."""
    if isinstance(a, (list, tuple)) and convert:
        a = np.asarray(a, dtype=np.float32)
    if isinstance(a, np.ndarray):
        if a.size != n:
            raise ValueError(f"{what}: expected {n} elements, got {a.size} (shape {a.shape})")
        if a.dtype != np.float32 or not a.flags.c_contiguous:
            if not convert:
                if a.dtype != np.float32:
                    raise TypeError(f"{what}: output buffer must be float32, got {a.dtype}")
                raise ValueError(f"{what}: output buffer must be C-contiguous")
            a = np.ascontiguousarray(a, dtype=np.float32)
        return C.c_void_p(a.ctypes.data), 0, a
    # torch tensor (duck-typed to keep torch optional at import time)
    if hasattr(a, "data_ptr"):
        if a.numel() != n:
            raise ValueError(f"{what}: expected {n} elements, got {a.numel()} (shape {tuple(a.shape)})")
        if str(a.dtype) != "torch.float32" or not a.is_contiguous():
            if not convert:
                if str(a.dtype) != "torch.float32":
                    raise TypeError(f"{what}: output buffer must be float32, got {a.dtype}")
                raise ValueError(f"{what}: output buffer must be contiguous")
            a = a.float().contiguous()
        return C.c_void_p(a.data_ptr()), (1 if a.is_cuda else 0), a
    raise TypeError(f"{what}: expected a numpy array or a torch tensor, got {type(a).__name__}")


def tissue_probe_input(sample, depth, maxt, channel, active=None):
    """in[n][5] of lr_kat_tissue / orc_kat_tissue: sample, tissue depth, maxt as floats; channel, active mask as uint32 bit patterns."""
    sample = np.asarray(sample, dtype=np.float32).reshape(-1); n = sample.shape[0]
    a = np.empty((n, 5), dtype=np.float32)
    a[:, 0] = sample; a[:, 1] = np.broadcast_to(np.asarray(depth, dtype=np.float32), (n,)); a[:, 2] = np.broadcast_to(np.asarray(maxt, dtype=np.float32), (n,))
    a[:, 3] = np.broadcast_to(np.asarray(channel, dtype=np.uint32), (n,)).view(np.float32) if np.ndim(channel) else np.full(n, channel, np.uint32).view(np.float32)
    act = np.ones(n, np.uint32) if active is None else np.broadcast_to(np.asarray(active, dtype=np.uint32), (n,)).copy()
    a[:, 4] = act.view(np.float32)
    return np.ascontiguousarray(a)


class Context:
    def __init__(self, device: int = 0, lib_path: str | None = None):
        self.lib = K.load_library(lib_path)
        self.ctx = C.c_void_p()
        st = self.lib.lr_create(int(device), C.byref(self.ctx))
        if st != K.LR_OK:
            K.check(self.lib, None, st, "lr_create")
        self.device = device
        self.scene: Scene | None = None

    def close(self):
        if getattr(self, "ctx", None):
            self.lib.lr_destroy(self.ctx); self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _chk(self, st, what):
        K.check(self.lib, self.ctx, st, what)

    # -- scene ---------------------------------------------------------------
    def upload(self, scene: Scene):
        pod = scene.pod()
        self._chk(self.lib.lr_scene_upload(self.ctx, C.byref(pod)), "lr_scene_upload")
        self.scene = scene; self._sensor_index = scene.sensor_index
        return self

    def set_stream(self, cuda_stream_ptr: int):
        self._chk(self.lib.lr_set_stream(self.ctx, C.c_void_p(cuda_stream_ptr)), "lr_set_stream")

    # -- forward ---------------------------------------------------------------
    def set_sensor(self, index: int):
        """lr_sensor_set: render the uploaded scene through another of its sensors (the `sensor` argument of mi.render)."""
        sc = self.scene
        if index != sc.sensor_index:
            sc.select_sensor(index)
        if index != self._sensor_index:
            self._chk(self.lib.lr_sensor_set(self.ctx, C.byref(sc.sensors[index])), "lr_sensor_set")
            self._sensor_index = index

    def render(self, seed: int = 0, spp: int | None = None, out=None, sensor: int | None = None):
        if sensor is not None:
            self.set_sensor(sensor)
        sc = self.scene; W, H = sc.film_size; spp = spp or sc.spp
        if out is None:
            out = np.empty((H, W, sc.out_channels), dtype=np.float32)
        p, dev, _ = _ptr(out, H * W * sc.out_channels, "render(out=)")
        self._chk(self.lib.lr_render(self.ctx, seed & 0xffffffff, spp, p, dev), "lr_render")
        return out

    def film_clear(self):
        self._chk(self.lib.lr_film_clear(self.ctx), "lr_film_clear")

    def render_lanes(self, seed, spp, begin, end):
        self._chk(self.lib.lr_render_lanes(self.ctx, seed & 0xffffffff, spp, begin, end), "lr_render_lanes")

    def set_lane_map(self, period: int = 0, offset: int = 0, length: int = 0):
        """lr_set_lane_map: the following render / sample / backward calls work on the lanes l with (l mod period) in
        [offset, offset + length) -- `multi.sample_share` gives the triple for rank r of N; length 0 removes the map."""
        self._chk(self.lib.lr_set_lane_map(self.ctx, period, offset, length), "lr_set_lane_map")

    def film_channels(self):
        return self.lib.lr_film_channels(self.ctx)

    def film_read(self, out=None):
        W, H = self.scene.film_size
        if out is None:
            out = np.empty((H, W, self.film_channels()), dtype=np.float32)
        p, dev, _ = _ptr(out, H * W * self.film_channels(), "film_read(out=)"); self._chk(self.lib.lr_film_read(self.ctx, p, dev), "lr_film_read"); return out

    def film_write(self, block):
        W, H = self.scene.film_size
        p, dev, _keep = _ptr(block, H * W * self.film_channels(), "film_write(block)", convert=True); self._chk(self.lib.lr_film_write(self.ctx, p, dev), "lr_film_write")

    def film_develop(self, out=None):
        sc = self.scene; W, H = sc.film_size
        if out is None:
            out = np.empty((H, W, sc.out_channels), dtype=np.float32)
        p, dev, _ = _ptr(out, H * W * sc.out_channels, "film_develop(out=)"); self._chk(self.lib.lr_film_develop(self.ctx, p, dev), "lr_film_develop"); return out

    def film_device_ptr(self):
        return self.lib.lr_film_device_ptr(self.ctx)

    def sample_lanes(self, seed, spp, begin, end, out=None):
        """Per-lane radiance [(end - begin) * n_passes, 4]; a render beyond 2^32 - 1 lanes is split into passes
        (src/render/integrator.cpp:275-293) and every lane then has one row per pass: row = (lane - begin) * n_passes + pass."""
        from .multi import lane_count
        W, H = self.scene.film_size; rows = (end - begin) * lane_count(W, H, spp)[2]
        if out is None:
            out = np.empty((rows, 4), dtype=np.float32)
        p, dev, _ = _ptr(out, rows * 4, "sample_lanes(out=)")
        self._chk(self.lib.lr_sample_lanes(self.ctx, seed & 0xffffffff, spp, begin, end, p, dev), "lr_sample_lanes")
        return out

    # -- adjoint -----------------------------------------------------------------
    def render_backward(self, grad_in, seed: int = 0, spp: int | None = None, begin: int = 0, end: int = 0):
        spp = spp or self.scene.spp
        W, H = self.scene.film_size
        p, dev, _keep = _ptr(grad_in, H * W * (self.film_channels() - 1), "render_backward(grad_in)", convert=True)
        self._chk(self.lib.lr_render_backward(self.ctx, seed & 0xffffffff, spp, p, dev, begin, end), "lr_render_backward")

    def grid_update(self, grid: int, data):
        p, dev, _keep = _ptr(data, self.lib.lr_grid_size(self.ctx, grid), "grid_update(data)", convert=True)
        self._chk(self.lib.lr_grid_update(self.ctx, grid, p, dev), "lr_grid_update")

    def grid_grad_zero(self, grid: int):
        self._chk(self.lib.lr_grid_grad_zero(self.ctx, grid), "lr_grid_grad_zero")

    def grid_grad_read(self, grid: int, out=None):
        n = self.lib.lr_grid_size(self.ctx, grid)
        if out is None:
            out = np.empty(n, dtype=np.float32)
        p, dev, _ = _ptr(out, n, "grid_grad_read(out=)"); self._chk(self.lib.lr_grid_grad_read(self.ctx, grid, p, dev), "lr_grid_grad_read"); return out

    def grid_grad_device_ptr(self, grid: int):
        return self.lib.lr_grid_grad_device_ptr(self.ctx, grid)

    def grid_size(self, grid: int):
        return self.lib.lr_grid_size(self.ctx, grid)

    # -- probes / control ------------------------------------------------------------
    def ray_intersect(self, rays: np.ndarray):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        hits = np.empty((rays.shape[0], 4), dtype=np.float32)
        self._chk(self.lib.lr_ray_intersect(self.ctx, C.c_void_p(rays.ctypes.data), rays.shape[0], C.c_void_p(hits.ctypes.data)), "lr_ray_intersect")
        return hits

    def ray_intersect_noalloc(self, rays, hits=None):
        """lr_ray_intersect_noalloc: the same query with its scratch in the library's static device array (no cudaMalloc; at
        most `lr_noalloc_max_rays()` rays per call).  numpy in -> numpy out; a CUDA torch tensor [n, 8] in -> the kernel is
        enqueued on the context's stream and writes the CUDA tensor `hits` [n, 4], which the caller provides."""
        if isinstance(rays, np.ndarray) or not hasattr(rays, "data_ptr"):
            rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
            hits = np.empty((rays.shape[0], 4), dtype=np.float32)
            self._chk(self.lib.lr_ray_intersect_noalloc(self.ctx, C.c_void_p(rays.ctypes.data), rays.shape[0], C.c_void_p(hits.ctypes.data), 0),
                      "lr_ray_intersect_noalloc")
            return hits
        if hits is None or not rays.is_cuda or not hits.is_cuda or not rays.is_contiguous() or not hits.is_contiguous():
            raise ValueError("device call: contiguous CUDA float32 tensors rays [n, 8] and hits [n, 4] are both required")
        self._chk(self.lib.lr_ray_intersect_noalloc(self.ctx, C.c_void_p(rays.data_ptr()), rays.shape[0], C.c_void_p(hits.data_ptr()), 1),
                  "lr_ray_intersect_noalloc")
        return hits

    # -- device known-answer probes (lr_kat_*): numpy in, numpy out ---------------------------------
    @staticmethod
    def _f32(a, cols):
        return np.ascontiguousarray(a, dtype=np.float32).reshape(-1, cols)

    def kat_bsdf_eval_pdf(self, bsdf, wi, wo):
        wi = self._f32(wi, 3); wo = self._f32(wo, 3); out = np.empty((wi.shape[0], 4), dtype=np.float32)
        self._chk(self.lib.lr_kat_bsdf_eval_pdf(self.ctx, C.byref(bsdf), C.c_void_p(wi.ctypes.data), C.c_void_p(wo.ctypes.data), wi.shape[0], C.c_void_p(out.ctypes.data)), "lr_kat_bsdf_eval_pdf")
        return out[:, :3], out[:, 3]

    def kat_bsdf_sample(self, bsdf, wi, u):
        wi = self._f32(wi, 3); u = self._f32(u, 3); out = np.empty((wi.shape[0], 9), dtype=np.float32)
        self._chk(self.lib.lr_kat_bsdf_sample(self.ctx, C.byref(bsdf), C.c_void_p(wi.ctypes.data), C.c_void_p(u.ctypes.data), wi.shape[0], C.c_void_p(out.ctypes.data)), "lr_kat_bsdf_sample")
        return {"wo": out[:, 0:3], "pdf": out[:, 3], "eta": out[:, 4], "weight": out[:, 5:8], "sampled_type": out[:, 8].copy().view(np.uint32)}

    def kat_phase(self, kind, g, wi, u):
        wi = self._f32(wi, 3); u = self._f32(u, 2); out = np.empty((wi.shape[0], 5), dtype=np.float32)
        self._chk(self.lib.lr_kat_phase(self.ctx, kind, g, C.c_void_p(wi.ctypes.data), C.c_void_p(u.ctypes.data), wi.shape[0], C.c_void_p(out.ctypes.data)), "lr_kat_phase")
        return out[:, :3], out[:, 3], out[:, 4]

    def kat_grid_eval(self, grid, p, use_fat=True):
        p = self._f32(p, 3); out = np.empty((p.shape[0], 3), dtype=np.float32)
        self._chk(self.lib.lr_kat_grid_eval(self.ctx, grid, C.c_void_p(p.ctypes.data), p.shape[0], C.c_void_p(out.ctypes.data), 1 if use_fat else 0), "lr_kat_grid_eval")
        return out

    def kat_emitter_sample(self, ref_p, u):
        a = np.concatenate([self._f32(ref_p, 3), self._f32(u, 2)], axis=1); a = np.ascontiguousarray(a); out = np.empty((a.shape[0], 7), dtype=np.float32)
        self._chk(self.lib.lr_kat_emitter_sample(self.ctx, C.c_void_p(a.ctypes.data), a.shape[0], C.c_void_p(out.ctypes.data)), "lr_kat_emitter_sample")
        return out[:, :3], out[:, 3], out[:, 4:7]

    def kat_env_eval_pdf(self, d):
        d = self._f32(d, 3); out = np.empty((d.shape[0], 4), dtype=np.float32)
        self._chk(self.lib.lr_kat_env_eval_pdf(self.ctx, C.c_void_p(d.ctypes.data), d.shape[0], C.c_void_p(out.ctypes.data)), "lr_kat_env_eval_pdf")
        return out[:, :3], out[:, 3]

    def kat_tissue(self, medium, sample, depth, maxt, channel, active=None):
        """Tissue media probe: (distance, element type, mei.t, transmittance, sigma_t[ch], combined_extinction[ch]) per query."""
        a = tissue_probe_input(sample, depth, maxt, channel, active); out = np.empty((a.shape[0], 8), dtype=np.float32)
        self._chk(self.lib.lr_kat_tissue(self.ctx, C.byref(medium), C.c_void_p(a.ctypes.data), a.shape[0], C.c_void_p(out.ctypes.data)), "lr_kat_tissue")
        return out

    def cancel(self):
        self.lib.lr_cancel(self.ctx)

    def profile(self, on: bool = True):
        """Per-stage timing mode of the wavefront volpath (one batch in flight, CUDA events around every stage launch)."""
        self._chk(self.lib.lr_profile_enable(self.ctx, 1 if on else 0), "lr_profile_enable")

    def stage_profile(self) -> dict:
        p = K.lr_stage_profile(); self._chk(self.lib.lr_profile_get(self.ctx, C.byref(p)), "lr_profile_get")
        return {self.lib.lr_stage_name(i).decode(): {"ms": p.ms[i], "launches": int(p.launches[i]), "items": int(p.items[i])} for i in range(p.n_stages)}

    def stats(self) -> dict:
        s = K.lr_stats(); self._chk(self.lib.lr_stats_get(self.ctx, C.byref(s)), "lr_stats_get")
        return {k: getattr(s, k) for k, _ in K.lr_stats._fields_}
