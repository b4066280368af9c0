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


# ---- the fork's own scenes: biovolpath + tissue media ------------------------------------------------
# property values of scenes/Liver-SingleMesh/mitsuba3/scene.xml:51-82 (the MultiMesh file carries the same numbers, :53-89)
_CAPSULE_PROPS = {
    "sigma_collagen1_R": 3.146124563777685, "sigma_collagen1_G": 2.2189004838302524, "sigma_collagen1_B": 1.5741115169422308,
    "sigma_collagen2_R": 2.6853117983771604, "sigma_collagen2_G": 1.8938982001080131, "sigma_collagen2_B": 1.3435514528168673,
    "sigma_collagen3_R": 0.0033151997510829138, "sigma_collagen3_G": 0.002338145926059276, "sigma_collagen3_B": 0.0016587054973047745,
    "sigma_collagen4_R": 0.023206398257580394, "sigma_collagen4_G": 0.01636702148241493, "sigma_collagen4_B": 0.01161093848113342,
    "sigma_elastin1_R": 0.1062817070956702, "sigma_elastin1_G": 0.29006947548901363, "sigma_elastin1_B": 0.7164288436612656,
    "sigma_elastin2_R": 0.3938675027663073, "sigma_elastin2_G": 1.074963350341639, "sigma_elastin2_B": 2.6550010088623375,
    "sigma_elastin3_R": 0.5293245804372595, "sigma_elastin3_G": 1.4446597406707737, "sigma_elastin3_B": 3.5680965939208136,
    "sigma_elastin4_R": 0.18130408857496683, "sigma_elastin4_G": 0.49482439936361144, "sigma_elastin4_B": 1.2221433215398063,
}
_PARENCHYMA_PROPS = {
    "sigma_blood": {"type": "rgb", "value": [0.004611074674964207, 0.20900034649954347, 0.24625187839886722]},
    "sigma_bile": {"type": "rgb", "value": [0.002160333333333333, 0.0030312499999999997, 0.025356499999999997]},
    "sigma_hepatocity": 269.26180490217416,
    "sigma_lipid_water": {"type": "rgb", "value": [0.004632281950333333, 0.00048109802439999993, 0.00106273247395]},
}


def _bump_dielectric():
    return {"type": "bumpmap", "scale": 0.005, "texture": {"type": "bitmap", "bitmap": tissue_texture(), "raw": True},
            "bsdf": {"type": "dielectric", "int_ior": 1.38, "ext_ior": 1.0}}


def liver_single_mesh_bio(res: int = 1080, spp: int = 256, max_depth: int = 12, sampler: str = "independent", surface: str = "bumpmap"):
    """Liver-SingleMesh as the fork ships it (scenes/Liver-SingleMesh/mitsuba3/scene.xml): `biovolpath`, one `liver` medium
    inside the bump-mapped dielectric organ, the cavity environment map.  Mesh, bump texture and environment map are this
    repository's synthetic stand-ins (the fork's assets are not redistributable here)."""
    mesh = _liver_mesh("liver2"); mesh["to_world"] = _LIVER_T
    mesh["bsdf"] = _bump_dielectric()
    if surface == "roughdielectric":      # not the fork's scene: a smooth lobe at the boundary, so that biovolpath's emitter sampling (and its shadow walk
        mesh["bsdf"] = {"type": "roughdielectric", "int_ior": 1.38, "ext_ior": 1.0, "alpha": 0.3, "distribution": "beckmann"}   # through tissue) is exercised
    mesh["interior"] = {"type": "liver", "scale": 1.0, "phase": {"type": "isotropic"}, **_CAPSULE_PROPS, **_PARENCHYMA_PROPS}
    sensor = _liver_sensor(res, spp); sensor["sampler"]["type"] = sampler
    return {"type": "scene", "integrator": {"type": "biovolpath", "max_depth": max_depth, "rr_depth": 5},
            "sensor": sensor, "liver": mesh, "env": _liver_env()}


def liver_multi_mesh_bio(res: int = 1080, spp: int = 256, max_depth: int = 12, sampler: str = "ldsampler"):
    """Liver-MultiMesh as its (partly commented-out) scene file lays it out: the capsule shell `liver1` filled with the
    `glissonCapsule` medium around the `liver2` organ filled with `parenchyma` (scene.xml:53-110), `ldsampler` (:18)."""
    outer = _liver_mesh("liver1"); outer["to_world"] = T().translate([-5.05, 5.05, -5.05]).scale(1.05)
    inner = _liver_mesh("liver2"); inner["to_world"] = _LIVER_T
    outer["bsdf"] = _bump_dielectric(); outer["interior"] = {"type": "ref", "id": "capsule"}
    inner["bsdf"] = {"type": "dielectric", "int_ior": 1.38, "ext_ior": 1.0}
    inner["interior"] = {"type": "ref", "id": "parenchyma"}; inner["exterior"] = {"type": "ref", "id": "capsule"}
    sensor = _liver_sensor(res, spp); sensor["sampler"]["type"] = sampler
    return {"type": "scene", "integrator": {"type": "biovolpath", "max_depth": max_depth, "rr_depth": 5},
            "sensor": sensor,
            "capsule": {"type": "glissonCapsule", "scale": 1.0, "phase": {"type": "isotropic"}, **_CAPSULE_PROPS},
            "parenchyma": {"type": "parenchyma", "scale": 1.0, "phase": {"type": "isotropic"}, **_PARENCHYMA_PROPS},
            "outer": outer, "inner": inner, "env": _liver_env()}
