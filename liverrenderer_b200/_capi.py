"""ctypes mirror of include/lr_api.h and loader of the CUDA backend library.

There is NO CPU fallback: if `liblr_b200.so` is missing or no CUDA device is
usable, every compute entry point raises.  (The CPU oracle under oracle/ is test
infrastructure and is never imported from this package.)
"""
from __future__ import annotations

import ctypes as C
import os

c_float_p = C.POINTER(C.c_float)
c_uint32_p = C.POINTER(C.c_uint32)

# ---- enums (values mirror lr_api.h) -----------------------------------------
LR_OK = 0
LR_API_VERSION = 4
LR_ERR = {-1: "invalid argument", -2: "no CUDA device (this backend has no CPU path)", -3: "CUDA error",
          -4: "unsupported plugin/option", -5: "no scene uploaded", -6: "cancelled", -7: "out of memory"}

INTEGRATOR_PATH, INTEGRATOR_VOLPATH, INTEGRATOR_PRBVOLPATH, INTEGRATOR_BIOVOLPATH, INTEGRATOR_VOLPATHMIS = 0, 1, 2, 3, 4
TRACKING_REFERENCE, TRACKING_DDA = 0, 1
SAMPLER_INDEPENDENT, SAMPLER_LD = 0, 1
BSDF_DIFFUSE, BSDF_ROUGHDIELECTRIC, BSDF_PRINCIPLED, BSDF_NULL, BSDF_DIELECTRIC = 0, 1, 2, 3, 4
MF_BECKMANN, MF_GGX = 0, 1
PHASE_ISOTROPIC, PHASE_HG = 0, 1
MEDIUM_HOMOGENEOUS, MEDIUM_HETEROGENEOUS, MEDIUM_LIVER, MEDIUM_GLISSON, MEDIUM_PARENCHYMA = 0, 1, 2, 3, 4
EMITTER_AREA, EMITTER_CONSTANT, EMITTER_ENVMAP = 0, 1, 2
RFILTER_BOX, RFILTER_GAUSSIAN, RFILTER_TENT = 0, 1, 2
FILTER_NEAREST, FILTER_TRILINEAR = 0, 1
WRAP_REPEAT, WRAP_MIRROR, WRAP_CLAMP = 0, 1, 2
TEX_BILINEAR, TEX_NEAREST = 0, 1
BUMP_FLIP_INVALID_NORMALS, BUMP_SHADOWING_FUNCTION = 1, 2

PR_HAS_ANISOTROPIC, PR_HAS_METALLIC, PR_HAS_SPEC_TRANS, PR_HAS_SPEC_TINT = 1, 2, 4, 8
PR_HAS_SHEEN, PR_HAS_SHEEN_TINT, PR_HAS_FLATNESS, PR_HAS_CLEARCOAT, PR_ETA_SPECULAR = 16, 32, 64, 128, 256


class lr_mesh(C.Structure):
    _fields_ = [("positions", c_float_p), ("normals", c_float_p), ("texcoords", c_float_p), ("indices", c_uint32_p),
                ("n_vertices", C.c_uint32), ("n_faces", C.c_uint32), ("bsdf", C.c_int32),
                ("interior_medium", C.c_int32), ("exterior_medium", C.c_int32), ("emitter", C.c_int32),
                ("flip_normals", C.c_uint32), ("reserved", C.c_uint32)]


class lr_bsdf(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("mf_kind", C.c_uint32), ("sample_visible", C.c_uint32), ("flags", C.c_uint32),
                ("reflectance", C.c_float * 3), ("transmittance", C.c_float * 3), ("eta", C.c_float),
                ("alpha_u", C.c_float), ("alpha_v", C.c_float),
                ("roughness", C.c_float), ("anisotropic", C.c_float), ("metallic", C.c_float), ("spec_trans", C.c_float),
                ("specular", C.c_float), ("spec_tint", C.c_float), ("sheen", C.c_float), ("sheen_tint", C.c_float),
                ("flatness", C.c_float), ("clearcoat", C.c_float), ("clearcoat_gloss", C.c_float),
                ("diff_refl_srate", C.c_float), ("main_spec_srate", C.c_float), ("clearcoat_srate", C.c_float),
                ("tex_kind", C.c_uint32), ("tex_color1", C.c_float * 3), ("tex_to_uv", C.c_float * 6),
                ("refl_tex", C.c_uint32), ("bump_tex", C.c_uint32), ("bump_scale", C.c_float), ("bump_flags", C.c_uint32)]


class lr_texture(C.Structure):
    _fields_ = [("data", c_float_p), ("width", C.c_uint32), ("height", C.c_uint32), ("channels", C.c_uint32),
                ("filter", C.c_uint32), ("wrap", C.c_uint32), ("reserved", C.c_uint32), ("to_uv", C.c_float * 6)]


class lr_grid(C.Structure):
    _fields_ = [("data", c_float_p), ("res", C.c_uint32 * 3), ("channels", C.c_uint32), ("to_local", C.c_float * 16),
                ("bbox_min", C.c_float * 3), ("bbox_max", C.c_float * 3), ("filter", C.c_uint32), ("wrap", C.c_uint32),
                ("max_value", C.c_float), ("differentiable", C.c_uint32)]


class lr_medium(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("phase", C.c_uint32), ("phase_g", C.c_float), ("scale", C.c_float),
                ("sigma_t_grid", C.c_int32), ("albedo_grid", C.c_int32), ("sigma_t_const", C.c_float * 3),
                ("albedo_const", C.c_float * 3), ("sample_emitters", C.c_uint32), ("has_spectral_extinction", C.c_uint32),
                # tissue media (liver / glissonCapsule / parenchyma), values as the plugins store them
                ("layer_limit", C.c_float * 4), ("sigma_collagen", (C.c_float * 3) * 4), ("sigma_elastin", (C.c_float * 3) * 4),
                ("sigma_blood", C.c_float * 3), ("sigma_bile", C.c_float * 3), ("sigma_lipid_water", C.c_float * 3),
                ("sigma_hepatocity", C.c_float)]


class lr_emitter(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("mesh", C.c_int32), ("radiance", C.c_float * 3), ("scale", C.c_float),
                ("is_rectangle", C.c_uint32), ("rect_origin", C.c_float * 3), ("rect_edge_u", C.c_float * 3),
                ("rect_edge_v", C.c_float * 3), ("rect_normal", C.c_float * 3), ("inv_area", C.c_float),
                ("pixels", c_float_p), ("width", C.c_uint32), ("height", C.c_uint32), ("to_world", C.c_float * 16),
                ("reserved", C.c_float * 2)]


class lr_sensor(C.Structure):
    _fields_ = [("to_world", C.c_float * 16), ("sample_to_camera", C.c_float * 16), ("near_clip", C.c_float),
                ("far_clip", C.c_float), ("film_width", C.c_uint32), ("film_height", C.c_uint32), ("rfilter", C.c_uint32),
                ("rfilter_radius", C.c_float), ("rfilter_stddev", C.c_float), ("has_alpha", C.c_uint32),
                ("medium", C.c_int32), ("crop_offset_x", C.c_uint32), ("crop_offset_y", C.c_uint32)]


class lr_integrator(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("max_depth", C.c_int32), ("rr_depth", C.c_int32), ("hide_emitters", C.c_uint32),
                ("tracking", C.c_uint32), ("base_seed", C.c_uint32), ("majorant_cell", C.c_uint32), ("sampler", C.c_uint32),
                ("spectral_mis", C.c_uint32)]


class lr_scene(C.Structure):
    _fields_ = [("api_version", C.c_uint32), ("n_meshes", C.c_uint32), ("n_bsdfs", C.c_uint32), ("n_media", C.c_uint32),
                ("n_grids", C.c_uint32), ("n_emitters", C.c_uint32), ("n_textures", C.c_uint32),
                ("textures", C.POINTER(lr_texture)), ("meshes", C.POINTER(lr_mesh)), ("bsdfs", C.POINTER(lr_bsdf)), ("media", C.POINTER(lr_medium)),
                ("grids", C.POINTER(lr_grid)), ("emitters", C.POINTER(lr_emitter)),
                ("sensor", lr_sensor), ("integrator", lr_integrator)]


class lr_stats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("rays", C.c_uint64), ("collisions", C.c_uint64), ("real_scatters", C.c_uint64),
                ("launches", C.c_uint64), ("device_ms", C.c_double), ("tracking_ms", C.c_double),
                ("tracking_launches", C.c_uint64)]


class lr_stage_profile(C.Structure):
    _fields_ = [("n_stages", C.c_uint32), ("reserved", C.c_uint32), ("ms", C.c_double * 16), ("launches", C.c_uint64 * 16), ("items", C.c_uint64 * 16)]


# every symbol include/lr_api.h declares (tests check the library exports all of them)
API_SYMBOLS = [
    "lr_api_version", "lr_create", "lr_destroy", "lr_last_error", "lr_scene_upload", "lr_render", "lr_film_clear",
    "lr_render_lanes", "lr_film_channels", "lr_film_read", "lr_film_write", "lr_film_develop", "lr_film_device_ptr",
    "lr_sample_lanes", "lr_render_backward", "lr_grid_update", "lr_grid_grad_zero", "lr_grid_grad_read",
    "lr_grid_grad_device_ptr", "lr_grid_size", "lr_set_lane_map", "lr_sensor_set", "lr_ray_intersect", "lr_ray_intersect_noalloc", "lr_noalloc_max_rays", "lr_cancel", "lr_stats_get", "lr_set_stream",
    "lr_profile_enable", "lr_profile_get", "lr_stage_name",
    "lr_kat_bsdf_eval_pdf", "lr_kat_bsdf_sample", "lr_kat_phase", "lr_kat_grid_eval", "lr_kat_emitter_sample", "lr_kat_env_eval_pdf",
    "lr_kat_tissue",
]

_LIB = None
LIB_PATH = os.environ.get("LR_B200_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "liblr_b200.so")   # LR_B200_LIB: A/B builds of the same backend


def load_library(path: str | None = None):
    """dlopen the CUDA backend and declare the prototypes.  Raises if it is missing."""
    global _LIB
    if _LIB is not None and path is None:
        return _LIB
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise RuntimeError(f"liverrenderer_b200: CUDA backend library not found at {p}; "
                           "run `python -c 'import __graft_entry__ as g; g.build()'` (there is no CPU fallback)")
    lib = C.CDLL(p)
    vp = C.c_void_p
    lib.lr_api_version.restype = C.c_int
    lib.lr_create.argtypes = [C.c_int, C.POINTER(vp)]; lib.lr_create.restype = C.c_int
    lib.lr_destroy.argtypes = [vp]; lib.lr_destroy.restype = None
    lib.lr_last_error.argtypes = [vp]; lib.lr_last_error.restype = C.c_char_p
    lib.lr_scene_upload.argtypes = [vp, C.POINTER(lr_scene)]; lib.lr_scene_upload.restype = C.c_int
    lib.lr_render.argtypes = [vp, C.c_uint32, C.c_uint32, vp, C.c_int]; lib.lr_render.restype = C.c_int
    lib.lr_film_clear.argtypes = [vp]; lib.lr_film_clear.restype = C.c_int
    lib.lr_render_lanes.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint64]; lib.lr_render_lanes.restype = C.c_int
    lib.lr_film_channels.argtypes = [vp]; lib.lr_film_channels.restype = C.c_int
    lib.lr_film_read.argtypes = [vp, vp, C.c_int]; lib.lr_film_read.restype = C.c_int
    lib.lr_film_write.argtypes = [vp, vp, C.c_int]; lib.lr_film_write.restype = C.c_int
    lib.lr_film_develop.argtypes = [vp, vp, C.c_int]; lib.lr_film_develop.restype = C.c_int
    lib.lr_film_device_ptr.argtypes = [vp]; lib.lr_film_device_ptr.restype = vp
    lib.lr_sample_lanes.argtypes = [vp, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint64, vp, C.c_int]; lib.lr_sample_lanes.restype = C.c_int
    lib.lr_render_backward.argtypes = [vp, C.c_uint32, C.c_uint32, vp, C.c_int, C.c_uint64, C.c_uint64]; lib.lr_render_backward.restype = C.c_int
    lib.lr_grid_update.argtypes = [vp, C.c_uint32, vp, C.c_int]; lib.lr_grid_update.restype = C.c_int
    lib.lr_grid_grad_zero.argtypes = [vp, C.c_uint32]; lib.lr_grid_grad_zero.restype = C.c_int
    lib.lr_grid_grad_read.argtypes = [vp, C.c_uint32, vp, C.c_int]; lib.lr_grid_grad_read.restype = C.c_int
    lib.lr_grid_grad_device_ptr.argtypes = [vp, C.c_uint32]; lib.lr_grid_grad_device_ptr.restype = vp
    lib.lr_grid_size.argtypes = [vp, C.c_uint32]; lib.lr_grid_size.restype = C.c_size_t
    lib.lr_ray_intersect.argtypes = [vp, vp, C.c_uint64, vp]; lib.lr_ray_intersect.restype = C.c_int
    lib.lr_ray_intersect_noalloc.argtypes = [vp, vp, C.c_uint64, vp, C.c_int]; lib.lr_ray_intersect_noalloc.restype = C.c_int
    lib.lr_noalloc_max_rays.argtypes = []; lib.lr_noalloc_max_rays.restype = C.c_uint64
    lib.lr_set_lane_map.argtypes = [vp, C.c_uint64, C.c_uint64, C.c_uint64]; lib.lr_set_lane_map.restype = C.c_int
    lib.lr_cancel.argtypes = [vp]; lib.lr_cancel.restype = C.c_int
    lib.lr_stats_get.argtypes = [vp, C.POINTER(lr_stats)]; lib.lr_stats_get.restype = C.c_int
    lib.lr_set_stream.argtypes = [vp, vp]; lib.lr_set_stream.restype = C.c_int
    lib.lr_profile_enable.argtypes = [vp, C.c_int]; lib.lr_profile_enable.restype = C.c_int
    lib.lr_profile_get.argtypes = [vp, C.POINTER(lr_stage_profile)]; lib.lr_profile_get.restype = C.c_int
    lib.lr_stage_name.argtypes = [C.c_uint32]; lib.lr_stage_name.restype = C.c_char_p
    lib.lr_kat_bsdf_eval_pdf.argtypes = [vp, C.POINTER(lr_bsdf), vp, vp, C.c_uint64, vp]; lib.lr_kat_bsdf_eval_pdf.restype = C.c_int
    lib.lr_kat_bsdf_sample.argtypes = [vp, C.POINTER(lr_bsdf), vp, vp, C.c_uint64, vp]; lib.lr_kat_bsdf_sample.restype = C.c_int
    lib.lr_kat_phase.argtypes = [vp, C.c_uint32, C.c_float, vp, vp, C.c_uint64, vp]; lib.lr_kat_phase.restype = C.c_int
    lib.lr_kat_grid_eval.argtypes = [vp, C.c_uint32, vp, C.c_uint64, vp, C.c_int]; lib.lr_kat_grid_eval.restype = C.c_int
    lib.lr_kat_emitter_sample.argtypes = [vp, vp, C.c_uint64, vp]; lib.lr_kat_emitter_sample.restype = C.c_int
    lib.lr_kat_env_eval_pdf.argtypes = [vp, vp, C.c_uint64, vp]; lib.lr_kat_env_eval_pdf.restype = C.c_int
    lib.lr_sensor_set.argtypes = [vp, C.POINTER(lr_sensor)]; lib.lr_sensor_set.restype = C.c_int
    lib.lr_kat_tissue.argtypes = [vp, C.POINTER(lr_medium), vp, C.c_uint64, vp]; lib.lr_kat_tissue.restype = C.c_int
    if path is None:
        _LIB = lib
    return lib


def check(lib, ctx, status: int, what: str):
    """Non-zero status -> RuntimeError, like Mitsuba's Throw() surfacing in Python
    (reference: include/mitsuba/core/logger.h:184-187)."""
    if status != LR_OK:
        msg = lib.lr_last_error(ctx)
        msg = msg.decode() if msg else LR_ERR.get(status, "unknown error")
        raise RuntimeError(f"{what}: {msg} (status {status})")
