"""Scene dictionaries of the reference's own participating-media render tests
(resources/data/tests/scenes/participating_media/*.xml + include/test-scene-base.xml),
rebuilt from the committed fixture tests/golden/participating_media.npz, plus the
reference's acceptance criterion (src/render/tests/test_renders.py:159-177, 203-228)."""
from __future__ import annotations

import os
import numpy as np

from liverrenderer_b200.transforms import T

_G = None


def golden():
    global _G
    if _G is None:
        _G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "participating_media.npz"))
    return _G


def _mesh(name, **kw):
    g = golden()
    d = {"type": "mesh", "positions": g[f"mesh_{name}_positions"], "indices": g[f"mesh_{name}_indices"],
         "normals": g[f"mesh_{name}_normals"], "texcoords": g[f"mesh_{name}_texcoords"]}
    d.update(kw)
    return d


def _base(integrator="volpath", res=48, spp=64, background=True, env_scale=1.0):
    # include/test-scene-base.xml: max_depth 6; include/test-scene-base-no-bg.xml: max_depth 16
    d = {"type": "scene",
         "integrator": {"type": integrator, "max_depth": 6 if background else 16},
         "sensor": {"type": "perspective", "near_clip": 1.0, "far_clip": 1000.0, # no fov => focal_length default 50mm (sensor.cpp:160-174)
                    "to_world": T().look_at(target=[0.0, -0.5, 0.0], origin=[0.0, 0.5, 4.0], up=[0.0, 1.0, 0.0]),
                    "film": {"type": "hdrfilm", "rfilter": {"type": "box"}, "width": res, "height": res,
                             "pixel_format": "rgb", "component_format": "float32"},
                    "sampler": {"type": "independent", "sample_count": spp}},
         "env": {"type": "envmap", "bitmap": golden()["museum"], "scale": env_scale}}
    if background:
        d["background"] = _mesh("background", bsdf={
            "type": "diffuse",
            "reflectance": {"type": "checkerboard", "color0": 0.2, "color1": 0.4,
                            "to_uv": T(np.diag([16.0, 16.0, 16.0, 1.0]))}})
    return d


# XML order scale -> translate: the first listed op is applied first => matrix = T * S
_VOL_T = T().translate([-0.5, -0.5, -0.5]).scale([1.5, 1.5, 1.5])


def scene(name: str, integrator: str = "volpath", res: int = 48, spp: int = 64):
    g = golden()
    if name == "test_hetero_sigmat":
        d = _base(integrator, res, spp)
        d["medium1"] = {"type": "heterogeneous", "scale": 5.0,
                        "sigma_t": {"type": "gridvolume", "grid": g["sigmat_vol"], "to_world": _VOL_T},
                        "albedo": {"type": "rgb", "value": [0.99, 0.5, 0.5]}, "phase": {"type": "isotropic"}}
        d["cube"] = _mesh("unit_cube_centered", bsdf={"type": "null"}, interior={"type": "ref", "id": "medium1"},
                          to_world=T().scale([1.5, 1.5, 1.5]))
        d["sphere"] = _mesh("sphere_base", bsdf={"type": "diffuse"})
    elif name == "test_hetero_sigmat_albedo":
        d = _base(integrator, res, spp)
        d["medium1"] = {"type": "heterogeneous", "scale": 5.0,
                        "sigma_t": {"type": "gridvolume", "grid": g["sigmat_vol"], "to_world": _VOL_T},
                        "albedo": {"type": "gridvolume", "grid": g["albedo_vol"], "to_world": _VOL_T},
                        "phase": {"type": "isotropic"}}
        d["cube"] = _mesh("unit_cube_centered", bsdf={"type": "null"}, interior={"type": "ref", "id": "medium1"},
                          to_world=T().scale([1.5, 1.5, 1.5]))
        d["sphere"] = _mesh("sphere_base", bsdf={"type": "diffuse"})
    elif name == "test_homo_null_envmap":
        d = _base(integrator, res, spp)
        d["medium1"] = {"type": "homogeneous", "scale": 3.0, "sigma_t": {"type": "rgb", "value": [1, 1, 1]},
                        "albedo": {"type": "rgb", "value": [0.99, 0.5, 0.5]}, "phase": {"type": "isotropic"}}
        d["tiny"] = _mesh("sphere_tiny", bsdf={"type": "null"}, interior={"type": "ref", "id": "medium1"})
        d["sphere"] = _mesh("sphere_base", bsdf={"type": "null"}, interior={"type": "ref", "id": "medium1"})
    elif name == "test_homo_dielectric_envmap":
        d = _base(integrator, res, spp, background=False)
        d["medium1"] = {"type": "homogeneous", "scale": 0.5, "sigma_t": {"type": "rgb", "value": [1, 1, 1]},
                        "albedo": {"type": "rgb", "value": [0.99, 0.7, 0.99]}, "phase": {"type": "isotropic"}}
        diel = {"type": "dielectric", "int_ior": "water", "ext_ior": "air"}
        d["tiny"] = _mesh("sphere_tiny", bsdf=dict(diel), interior={"type": "ref", "id": "medium1"})
        d["sphere"] = _mesh("sphere_base", bsdf=dict(diel), interior={"type": "ref", "id": "medium1"})
    else:
        raise KeyError(name)
    return d


GOLDEN_SCENES = ["test_hetero_sigmat", "test_hetero_sigmat_albedo", "test_homo_null_envmap", "test_homo_dielectric_envmap"]


# linear sRGB -> CIE XYZ (include/mitsuba/core/spectrum.h:398-400, src/core/bitmap.cpp:498-506)
RGB2XYZ = np.array([[0.412453, 0.357580, 0.180423], [0.212671, 0.715160, 0.072169], [0.019334, 0.119193, 0.950227]])


def to_xyz(img):
    return np.asarray(img, dtype=np.float64) @ RGB2XYZ.T


def z_test(mean, sample_count, reference, reference_var):
    """src/render/tests/test_renders.py:158-175 (variance floored at 1e-4)."""
    from scipy.special import erf
    reference_var = np.maximum(reference_var, 1e-4)
    z_stat = np.abs(mean - reference) * np.sqrt(sample_count / reference_var)
    cdf = (1.0 - erf(-z_stat / np.sqrt(2.0))) * 0.5
    return 2.0 * (1.0 - cdf)


def reference_accepts(img, name, sample_count, significance_level=0.01):
    """test_render acceptance (test_renders.py:181-228).  The reference compares in CIE XYZ: the
    `moment` integrator's AOVs are XYZ, the stored EXRs are xyz_to_rgb(mean) and xyz_to_rgb(variance)
    (:316-321) and are converted back with read_rgb_bmp_to_xyz (:124-126).  Sidak-corrected alpha
    with pixel_count = W*H; pass when count_nonzero(success)/3 >= 0.9975 * pixel_count.
    Returns (fraction of passing entries, pass/fail)."""
    g = golden(); ref = to_xyz(g[f"{name}_ref"]); var = to_xyz(g[f"{name}_var"])
    p = z_test(to_xyz(img), sample_count, ref, var)
    pixel_count = ref.shape[0] * ref.shape[1]
    alpha = 1.0 - (1.0 - significance_level) ** (1.0 / pixel_count)
    success = p > alpha
    ok = (np.count_nonzero(success) / 3) >= (0.9975 * pixel_count)
    return float(np.count_nonzero(success)) / float(success.size), bool(ok)
