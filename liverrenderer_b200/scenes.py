"""Scene dictionaries of the BASELINE.json configurations.

C1 is `mi.cornell_box()` itself; C2 follows MitsubaRunner.py:23-46,67-78; C3-C5 are the
`heterogeneous`+`gridvolume` restatements of the fork's liver scenes (the fork's own XMLs use
its custom `liver`/`biovolpath` plugins, which are outside this hot path).  Geometry comes from
the committed fixture liverrenderer_b200/data/liver_meshes.npz (made by tools/make_fixtures.py
from scenes/Liver-SingleMesh/mitsuba3/liver{1,2}.obj); the environment map is the 4x
box-downsampled fp16 copy of cavidade_latitude.exr, bilinearly upsampled back to 1024x512.
"""
from __future__ import annotations

import os
import numpy as np

from .transforms import T

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")


def sigmat_field(res: int) -> np.ndarray:
    """Closed-form density of the reference's generator
    (resources/data/tests/scenes/participating_media/create_volume_data.py:33-36),
    evaluated at (i+0)/res per axis; max 5.0.  Returns [z][y][x][1] fp32."""
    c = np.arange(res, dtype=np.float64) / res
    y, x = np.meshgrid(c, c, indexing="ij")                      # the field depends on z only through the block below
    f = np.maximum(0.0, (np.sin(16 * (x + 4 * ((y - 0.5) ** 2 - 0.25)) + 1) * 0.5)).astype(np.float32)
    blk = (x > 0.2) & (y > 0.7)
    out = np.empty((res, res, res, 1), dtype=np.float32); out[..., 0] = f[None]
    out[: int(np.count_nonzero(c < 0.5)), :, :, 0][:, blk] = 5.0
    return out


def tissue_texture(res: int = 512, seed: int = 7) -> np.ndarray:
    """Synthetic stand-in for the fork's `tissue_n.png` bump texture (scenes/Liver-SingleMesh/mitsuba3/scene.xml:31-35):
    a tileable band-limited RGB field in [0, 1], [res][res][3] fp32 (used `raw`)."""
    rng = np.random.default_rng(seed); c = np.arange(res, dtype=np.float64) / res
    y, x = np.meshgrid(c, c, indexing="ij"); out = np.zeros((res, res, 3))
    for ch in range(3):
        f = np.zeros((res, res))
        for _ in range(24):
            kx, ky = rng.integers(1, 24, size=2) * rng.choice([-1, 1], size=2); amp = 1.0 / np.hypot(kx, ky)
            f += amp * np.sin(2 * np.pi * (kx * x + ky * y) + rng.uniform(0, 2 * np.pi))
        out[..., ch] = (f - f.min()) / (f.max() - f.min())
    return np.ascontiguousarray(out, dtype=np.float32)


def _liver_mesh(name: str):
    z = np.load(os.path.join(_DATA, "liver_meshes.npz"))
    return {"type": "mesh", "positions": z[f"{name}_positions"], "indices": z[f"{name}_indices"],
            "normals": z[f"{name}_normals"], "texcoords": z[f"{name}_texcoords"]}


def cavity_envmap(width: int = 1024, height: int = 512) -> np.ndarray:
    small = np.load(os.path.join(_DATA, "cavity_env_256x128.npz"))["rgb"].astype(np.float32)
    h, w, _ = small.shape
    ys = (np.arange(height) + 0.5) * h / height - 0.5; xs = (np.arange(width) + 0.5) * w / width - 0.5
    y0 = np.clip(np.floor(ys).astype(int), 0, h - 1); y1 = np.clip(y0 + 1, 0, h - 1); wy = np.clip(ys - np.floor(ys), 0, 1)[:, None, None]
    x0 = np.floor(xs).astype(int) % w; x1 = (x0 + 1) % w; wx = (xs - np.floor(xs))[None, :, None]
    top = small[y0][:, x0] * (1 - wx) + small[y0][:, x1] * wx
    bot = small[y1][:, x0] * (1 - wx) + small[y1][:, x1] * wx
    return np.ascontiguousarray(top * (1 - wy) + bot * wy, dtype=np.float32)


def cornell_box_path(res: int = 256, spp: int = 64):
    """C1: mi.cornell_box(), 256x256, path max_depth 8, 64 spp."""
    from . import cornell_box
    d = cornell_box()
    d["sensor"]["film"]["width"] = res; d["sensor"]["film"]["height"] = res
    d["sensor"]["sampler"]["sample_count"] = spp
    return d


def cornell_box_fog(res: int = 1080, spp: int = 256):
    """C2: MitsubaRunner.py:23-46,67-78 -- homogeneous fog filling the box, sensor inside the medium."""
    d = cornell_box_path(res, spp)
    d["fog"] = {"type": "homogeneous", "sigma_t": {"type": "rgb", "value": [0.2, 0.2, 0.2]},
                "albedo": {"type": "rgb", "value": [0.75, 0.75, 0.75]}, "scale": 2.5, "phase": {"type": "isotropic"}}
    d["sensor"]["medium"] = {"type": "ref", "id": "fog"}
    d["integrator"] = {"type": "volpath", "max_depth": -1, "rr_depth": 5}
    return d


_LIVER_T = T().translate([-5, 5, -5])
_LIVER_BBOX = (np.array([-59.14, -32.92, -61.76]), np.array([-27.58, -2.78, -23.66]))


def _grid_to_world(bmin, bmax):
    return T().translate(bmin).scale(np.asarray(bmax) - np.asarray(bmin))


def _liver_sensor(res, spp, rgba=True):
    return {"type": "perspective", "fov": 45.0, "near_clip": 0.01, "far_clip": 10000.0,
            "to_world": T().look_at(origin=[7.481132, 5.343665, 6.507640], target=[6.826270, 5.011984, 5.896974],
                                    up=[-0.317370, 0.895343, -0.312469]),
            "sampler": {"type": "independent", "sample_count": spp},
            "film": {"type": "hdrfilm", "width": res, "height": res, "rfilter": {"type": "box"},
                     "pixel_format": "rgba" if rgba else "rgb", "component_format": "float32"}}


def _liver_env():
    return {"type": "envmap", "bitmap": cavity_envmap(), "scale": 2.5,
            # XML order rotate -> scale -> translate (scene.xml:93-101): first listed applied first => T * S * R
            "to_world": T().translate([-3, 3, 4]).scale(1).rotate([0.57735, 0.57735, 0.57735], 180)}


def liver_single_mesh(res: int = 1080, spp: int = 1024, grid_res: int = 256, phase: str = "isotropic",
                      tracking: str = "reference", max_depth: int = 12, surface: str = "roughdielectric"):
    """C3: Liver-SingleMesh restated with volpath + heterogeneous gridvolume.
    `surface="bumpmap"` keeps the fork's own boundary instead of the SURVEY substitute: `bumpmap` (scale 0.005,
    bitmap texture) over a smooth `dielectric` 1.38 (scene.xml:31-40), with a synthetic tissue texture."""
    mesh = _liver_mesh("liver2"); mesh["to_world"] = _LIVER_T
    pos = _LIVER_T.transform_affine(mesh["positions"].astype(np.float64))
    bmin, bmax = pos.min(0), pos.max(0)
    mesh["bsdf"] = {"type": "roughdielectric", "int_ior": 1.38, "ext_ior": 1.0, "alpha": 0.1, "distribution": "beckmann"}
    if surface == "bumpmap":
        mesh["bsdf"] = {"type": "bumpmap", "scale": 0.005,
                        "texture": {"type": "bitmap", "bitmap": tissue_texture(), "raw": True},
                        "bsdf": {"type": "dielectric", "int_ior": 1.38, "ext_ior": 1.0}}
    elif surface != "roughdielectric":
        raise ValueError(f"unknown surface '{surface}'")
    mesh["interior"] = {"type": "heterogeneous", "scale": 0.4,
                        "sigma_t": {"type": "gridvolume", "grid": sigmat_field(grid_res), "to_world": _grid_to_world(bmin, bmax)},
                        "albedo": {"type": "rgb", "value": [0.92, 0.45, 0.35]},
                        "phase": {"type": "hg", "g": 0.8} if phase == "hg" else {"type": "isotropic"}}
    return {"type": "scene", "integrator": {"type": "volpath", "max_depth": max_depth, "rr_depth": 5, "tracking": tracking},
            "sensor": _liver_sensor(res, spp), "liver": mesh, "env": _liver_env()}


def liver_multi_mesh(res: int = 1080, spp: int = 4096, tracking: str = "reference"):
    """C4: Liver-MultiMesh restated: capsule shell (principled) + parenchyma (roughdielectric), two media."""
    outer = _liver_mesh("liver1"); outer["to_world"] = T().translate([-5.05, 5.05, -5.05]).scale(1.05)
    inner = _liver_mesh("liver2"); inner["to_world"] = _LIVER_T
    po = outer["to_world"].transform_affine(outer["positions"].astype(np.float64))
    pi = _LIVER_T.transform_affine(inner["positions"].astype(np.float64))
    capsule = {"type": "heterogeneous", "scale": 1.2,
               "sigma_t": {"type": "gridvolume", "grid": sigmat_field(128), "to_world": _grid_to_world(po.min(0), po.max(0))},
               "albedo": {"type": "rgb", "value": [0.95, 0.9, 0.85]}, "phase": {"type": "hg", "g": 0.9}}
    paren = {"type": "heterogeneous", "scale": 0.4,
             "sigma_t": {"type": "gridvolume", "grid": sigmat_field(256), "to_world": _grid_to_world(pi.min(0), pi.max(0))},
             "albedo": {"type": "rgb", "value": [0.92, 0.45, 0.35]}, "phase": {"type": "isotropic"}}
    outer["bsdf"] = {"type": "principled", "base_color": {"type": "rgb", "value": [0.8, 0.3, 0.25]}, "roughness": 0.3,
                     "spec_trans": 0.9, "eta": 1.38}
    outer["interior"] = {"type": "ref", "id": "capsule"}
    inner["bsdf"] = {"type": "roughdielectric", "int_ior": 1.40, "ext_ior": 1.38, "alpha": 0.2}
    inner["interior"] = {"type": "ref", "id": "parenchyma"}; inner["exterior"] = {"type": "ref", "id": "capsule"}
    return {"type": "scene", "integrator": {"type": "volpath", "max_depth": 12, "rr_depth": 5, "tracking": tracking},
            "sensor": _liver_sensor(res, spp), "capsule": capsule, "parenchyma": paren,
            "outer": outer, "inner": inner, "env": _liver_env()}


def sphere_liver_inverse(res: int = 512, spp: int = 64, grid_res: int = 128, density=None, tracking: str = "reference"):
    """C5: SphereLiverConstEnv restated for prbvolpath (tutorials/inverse_rendering/volume_optimization.ipynb)."""
    if density is None:
        density = 0.25 * sigmat_field(grid_res) / 5.0 + 0.002
    sphere = {"type": "sphere", "subdivisions": 5,
              "bsdf": {"type": "roughdielectric", "int_ior": 1.5, "ext_ior": 1.0, "alpha": 0.1},
              "interior": {"type": "heterogeneous", "scale": 40.0,
                           "sigma_t": {"type": "gridvolume", "grid": np.asarray(density, dtype=np.float32),
                                       "to_world": T().translate([-1, -1, -1]).scale(2.0)},
                           "albedo": {"type": "rgb", "value": [0.8, 0.8, 0.8]}, "phase": {"type": "isotropic"}}}
    floor = {"type": "rectangle", "to_world": T().translate([0, -1, 0]).scale([25, 1, 25]).rotate([1, 0, 0], -90),
             "bsdf": {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.5, 0.5, 0.5]}}}
    return {"type": "scene", "integrator": {"type": "prbvolpath", "max_depth": 8, "rr_depth": 5, "tracking": tracking},
            "sensor": {"type": "perspective", "fov": 35.0, "near_clip": 0.01, "far_clip": 1000.0,
                       "to_world": T().look_at(origin=[0, 5, -10], target=[0, 0, 0], up=[0, 1, 0]),
                       "sampler": {"type": "independent", "sample_count": spp},
                       "film": {"type": "hdrfilm", "width": res, "height": res, "rfilter": {"type": "box"},
                                "pixel_format": "rgb", "component_format": "float32"}},
            "sphere": sphere, "floor": floor, "env": {"type": "constant", "radiance": {"type": "rgb", "value": [1, 1, 1]}}}
