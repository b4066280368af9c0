"""ctypes binding of the CPU oracle (oracle/liborc.so).  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs, never by the product package."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import numpy as np

from liverrenderer_b200 import _capi as K

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "liborc.so")
LIB_FAST = os.path.join(ROOT, "oracle", "_fast", "liborc_fast.so")
FAST_FLAGS = "-O3 -march=native"            # the TIMING build (bench.py's CPU legs only); parity always uses liborc.so (-O2 -ffp-contract=off)
_lib = None
_libs = {}


def build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])


def build_fast():
    """Compiles the oracle once more with FAST_FLAGS ON THE MACHINE IT WILL BE TIMED ON (-march=native is not portable, so this
    binary never travels: it is rebuilt, ~10 s, whenever the CPU differs from the one recorded in the stamp)."""
    d = os.path.dirname(LIB_FAST); os.makedirs(d, exist_ok=True)
    try:
        cpu = [l for l in open("/proc/cpuinfo") if l.startswith(("model name", "flags"))][:2]
    except OSError:
        cpu = []
    stamp = FAST_FLAGS + "\n" + "".join(cpu); sp = os.path.join(d, "stamp")
    src = os.path.join(ROOT, "oracle")
    newest = max(os.path.getmtime(os.path.join(src, f)) for f in os.listdir(src) if f.endswith((".cpp", ".h", ".inl")))
    if not (os.path.exists(LIB_FAST) and os.path.exists(sp) and open(sp).read() == stamp and os.path.getmtime(LIB_FAST) >= newest):
        subprocess.check_call(["make", "-s", "-B", "-C", src, "fast", "FASTFLAGS=" + FAST_FLAGS])
        open(sp, "w").write(stamp)
    return LIB_FAST


def lib(fast: bool = False):
    global _lib
    if fast:
        if "fast" not in _libs:
            _libs["fast"] = _bind(C.CDLL(build_fast()))
        return _libs["fast"]
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB):
        build()
    _lib = _bind(C.CDLL(LIB))
    return _lib


def _bind(L):
    vp = C.c_void_p
    L.orc_last_error.restype = C.c_char_p
    L.orc_scene_create.argtypes = [C.POINTER(K.lr_scene)]; L.orc_scene_create.restype = vp
    L.orc_scene_destroy.argtypes = [vp]
    L.orc_sample_lanes.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint64, vp, C.c_int]
    L.orc_film_clear.argtypes = [vp]
    L.orc_render_lanes.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint64, C.c_int]
    L.orc_film_channels.argtypes = [vp]; L.orc_film_channels.restype = C.c_int
    L.orc_film_read.argtypes = [vp, vp]; L.orc_film_develop.argtypes = [vp, vp]
    L.orc_render.argtypes = [vp, C.c_uint32, C.c_uint32, vp, C.c_int]
    L.orc_render_backward.argtypes = [vp, C.c_uint32, C.c_uint32, vp, C.c_uint64, C.c_uint64, C.c_int]
    L.orc_grid_size.argtypes = [vp, C.c_uint32]; L.orc_grid_size.restype = C.c_size_t
    L.orc_grid_grad_zero.argtypes = [vp, C.c_uint32]; L.orc_grid_grad_read.argtypes = [vp, C.c_uint32, vp]
    L.orc_grid_update.argtypes = [vp, C.c_uint32, vp]
    L.orc_ray_intersect.argtypes = [vp, vp, C.c_uint64, vp]
    L.orc_stats.argtypes = [vp, vp]; L.orc_stats_reset.argtypes = [vp]
    L.orc_kat_tea32.argtypes = [C.c_uint32, C.c_uint32, C.c_int, vp]
    L.orc_kat_tea_float32.argtypes = [C.c_uint32, C.c_uint32, C.c_int]; L.orc_kat_tea_float32.restype = C.c_float
    L.orc_kat_pcg32_floats.argtypes = [C.c_uint64, C.c_uint64, C.c_int, vp]
    L.orc_kat_pcg32_uints.argtypes = [C.c_uint64, C.c_uint64, C.c_int, vp]
    L.orc_kat_sampler_floats.argtypes = [C.c_uint32, C.c_uint32, C.c_int, vp]
    L.orc_kat_ldsampler.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int, vp]
    L.orc_kat_bsdf_eval_pdf.argtypes = [C.POINTER(K.lr_bsdf), vp, vp, vp, vp]
    L.orc_kat_bsdf_sample.argtypes = [C.POINTER(K.lr_bsdf), vp, C.c_float, C.c_float, C.c_float, vp]; L.orc_kat_bsdf_sample.restype = C.c_uint32
    L.orc_kat_texture.argtypes = [C.POINTER(K.lr_texture), C.c_int, vp, vp]
    L.orc_kat_bump_frame.argtypes = [C.POINTER(K.lr_bsdf), C.POINTER(K.lr_texture), vp, vp]
    L.orc_kat_phase_sample.argtypes = [C.c_uint32, C.c_float, vp, C.c_float, C.c_float, vp]
    L.orc_kat_phase_eval.argtypes = [C.c_uint32, C.c_float, vp, vp]; L.orc_kat_phase_eval.restype = C.c_float
    L.orc_kat_grid_eval.argtypes = [vp, C.c_uint32, vp, vp]
    L.orc_kat_rfilter_eval.argtypes = [vp, C.c_float]; L.orc_kat_rfilter_eval.restype = C.c_float
    L.orc_kat_camera_ray.argtypes = [vp, C.c_float, C.c_float, vp]
    L.orc_kat_env_eval.argtypes = [vp, vp, vp]; L.orc_kat_env_sample.argtypes = [vp, vp, C.c_float, C.c_float, vp]
    L.orc_kat_env_pdf.argtypes = [vp, vp]; L.orc_kat_env_pdf.restype = C.c_float
    L.orc_kat_erfinv.argtypes = [C.c_float]; L.orc_kat_erfinv.restype = C.c_float
    L.orc_kat_tissue.argtypes = [C.POINTER(K.lr_medium), vp, C.c_uint64, vp]
    return L


def _p(a):
    return C.c_void_p(a.ctypes.data)


def f3(v):
    return np.ascontiguousarray(v, dtype=np.float32)


def texture_probe(tex, uv):
    """eval rgb(3), eval_1, eval_1_grad(2) of lr_texture `tex` at each uv pair -> [n, 6]."""
    uv = np.ascontiguousarray(uv, dtype=np.float32).reshape(-1, 2); out = np.empty((uv.shape[0], 6), dtype=np.float32)
    lib().orc_kat_texture(C.byref(tex), uv.shape[0], _p(uv), _p(out)); return out


def bump_frame(bsdf, tex, n, sh_n, sh_s, dp_du, dp_dv, wi, uv):
    """BumpMap::frame for a synthetic interaction -> rows n, s, t of the perturbed frame."""
    si = np.concatenate([n, sh_n, sh_s, dp_du, dp_dv, wi, uv]).astype(np.float32); out = np.empty(9, dtype=np.float32)
    lib().orc_kat_bump_frame(C.byref(bsdf), C.byref(tex), _p(si), _p(out)); return out.reshape(3, 3)


def tissue_probe(medium, sample, depth, maxt, channel, active=None):
    """orc_kat_tissue: same in/out layout as Backend.kat_tissue."""
    from liverrenderer_b200.backend import tissue_probe_input
    a = tissue_probe_input(sample, depth, maxt, channel, active); out = np.empty((a.shape[0], 8), dtype=np.float32)
    lib().orc_kat_tissue(C.byref(medium), _p(a), a.shape[0], _p(out))
    return out


class OracleScene:
    """CPU-oracle twin of liverrenderer_b200.Context for one scene."""

    def __init__(self, scene, threads: int = 0, fast: bool = False):
        """`fast`: the -O3 -march=native timing build (bench.py); parity tests never set it."""
        self.L = lib(fast); self.scene = scene; self.threads = threads
        pod = scene.pod()
        self.h = self.L.orc_scene_create(C.byref(pod))
        if not self.h:
            raise RuntimeError(self.L.orc_last_error().decode())

    def __del__(self):
        try:
            if self.h:
                self.L.orc_scene_destroy(self.h); self.h = None
        except Exception:
            pass

    def sample_lanes(self, seed, spp, begin, end):
        out = np.empty((end - begin, 4), dtype=np.float32)
        self.L.orc_sample_lanes(self.h, seed, spp, begin, end, _p(out), self.threads); return out

    def render(self, seed=0, spp=None):
        sc = self.scene; W, H = sc.film_size; spp = spp or sc.spp
        out = np.empty((H, W, sc.out_channels), dtype=np.float32)
        self.L.orc_render(self.h, seed, spp, _p(out), self.threads); return out

    def film_clear(self):
        self.L.orc_film_clear(self.h)

    def render_lanes(self, seed, spp, begin, end):
        self.L.orc_render_lanes(self.h, seed, spp, begin, end, self.threads)

    def film_read(self):
        W, H = self.scene.film_size
        out = np.empty((H, W, self.L.orc_film_channels(self.h)), dtype=np.float32)
        self.L.orc_film_read(self.h, _p(out)); return out

    def film_develop(self):
        sc = self.scene; W, H = sc.film_size
        out = np.empty((H, W, sc.out_channels), dtype=np.float32)
        self.L.orc_film_develop(self.h, _p(out)); return out

    def render_backward(self, grad_in, seed=0, spp=None, begin=0, end=0):
        g = np.ascontiguousarray(grad_in, dtype=np.float32)
        self.L.orc_render_backward(self.h, seed, spp or self.scene.spp, _p(g), begin, end, self.threads)

    def grid_grad_zero(self, g):
        self.L.orc_grid_grad_zero(self.h, g)

    def grid_grad_read(self, g):
        out = np.empty(self.L.orc_grid_size(self.h, g), dtype=np.float32)
        self.L.orc_grid_grad_read(self.h, g, _p(out)); return out

    def grid_update(self, g, data):
        d = np.ascontiguousarray(data, dtype=np.float32); self.L.orc_grid_update(self.h, g, _p(d))

    def ray_intersect(self, rays):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        hits = np.empty((rays.shape[0], 4), dtype=np.float32)
        self.L.orc_ray_intersect(self.h, _p(rays), rays.shape[0], _p(hits)); return hits

    def stats(self):
        a = np.zeros(4, dtype=np.uint64); self.L.orc_stats(self.h, _p(a))
        return dict(samples=int(a[0]), rays=int(a[1]), collisions=int(a[2]), real_scatters=int(a[3]))

    def stats_reset(self):
        self.L.orc_stats_reset(self.h)
