"""Scene description -> POD `lr_scene` (the host half of the drop-in boundary).

Mirrors the subset of `mi.load_dict` (src/core/parser.cpp + plugin constructors)
that the hot path needs: plugin names, property names and defaults follow the
reference plugins cited next to each builder.  The result owns every numpy
buffer the POD points into.
"""
from __future__ import annotations

import ctypes as C
import os
import numpy as np

from . import _capi as K
from . import shapes as S
from .transforms import ScalarTransform4f, parse_fov, perspective_projection
from .volume_io import read_vol
from .image_io import read_texture, texture_from_pixels


_TISSUE_MEDIA = {"liver": K.MEDIUM_LIVER, "glissonCapsule": K.MEDIUM_GLISSON, "parenchyma": K.MEDIUM_PARENCHYMA}


def _rgb(v, default=None):
    if v is None:
        v = default
    if isinstance(v, dict):
        t = v.get("type", "rgb")
        if t in ("rgb", "srgb", "uniform"):
            v = v["value"]
        else:
            raise NotImplementedError(f"texture/spectrum type '{t}' is outside the hot path (constant rgb only)")
    a = np.asarray(v, dtype=np.float32).reshape(-1)
    if a.size == 1:
        a = np.repeat(a, 3)
    if a.size != 3:
        raise ValueError("expected a float or an RGB triple")
    return a


def _flt(v, default=None):
    if v is None:
        v = default
    if isinstance(v, dict):
        v = v["value"]
    return float(np.asarray(v).reshape(-1)[0])


def _f16(m):
    return (C.c_float * 16)(*np.asarray(m, dtype=np.float32).reshape(-1))


def _f3(v):
    return (C.c_float * 3)(*np.asarray(v, dtype=np.float32).reshape(-1)[:3])


def _xf(d, key="to_world"):
    t = d.get(key)
    if t is None:
        return ScalarTransform4f()
    if isinstance(t, ScalarTransform4f):
        return t
    return ScalarTransform4f(np.asarray(t, dtype=np.float64))


def _uv6(d, key="to_uv"):
    """`to_uv` as a 2x3 affine: a 3x3 (ScalarTransform3f) or a 4x4 (ScalarTransform4f, z ignored) matrix."""
    t = d.get(key)
    if t is not None and not isinstance(t, ScalarTransform4f) and np.asarray(t).shape == (3, 3):
        m = np.asarray(t, dtype=np.float64)
        return (C.c_float * 6)(m[0, 0], m[0, 1], m[0, 2], m[1, 0], m[1, 1], m[1, 2])
    m = _xf(d, key).matrix
    return (C.c_float * 6)(m[0, 0], m[0, 1], m[0, 3], m[1, 0], m[1, 1], m[1, 3])


_IOR = {"vacuum": 1.0, "air": 1.000277, "water": 1.3330, "bk7": 1.5046, "diamond": 2.419, "glass": 1.5046}


def _ior(v, default):
    if v is None:
        v = default
    if isinstance(v, str):
        return _IOR[v.lower()]
    return float(v)


def ld_sample_count(spp: int) -> int:
    """LowDiscrepancySampler::set_sample_count (src/samplers/ldsampler.cpp:80-89): the sample count is rounded up to a
    square power of two (4, 16, 64, 256, 1024, ...)."""
    res = 2
    while res * res < spp:
        res += 1
        res = 1 << (res - 1).bit_length()          # math::round_to_power_of_two
    return res * res


class Scene:
    """Flattened scene + the context handle of the backend it was uploaded to."""

    def __init__(self, d: dict, base_dir: str = "."):
        if d.get("type", "scene") != "scene":
            raise ValueError("load_dict(): top-level dictionary must be of type 'scene'")
        self.base_dir = base_dir
        self._keep = []                       # numpy buffers referenced by the POD
        self.meshes, self.bsdfs, self.media, self.grids, self.emitters = [], [], [], [], []
        self.textures = []
        self._bsdf_ids, self._medium_ids = {}, {}
        self.grid_arrays = []                 # numpy views of grid data, index-aligned with self.grids
        self.param_keys = {}                  # 'shape.interior_medium.sigma_t.data' -> grid index
        self.sensor = None; self.integrator = None
        self.spp = 4
        self._named = {k: v for k, v in d.items() if isinstance(v, dict)}
        self._integrator_type = next((v.get("type") for v in self._named.values() if v.get("type") in ("path", "volpath", "prbvolpath", "biovolpath", "volpathmis")), "path")
        # first pass: named BSDFs / media so that {'type':'ref','id':...} resolves
        for name, v in self._named.items():
            t = v.get("type")
            if t in ("diffuse", "roughdielectric", "principled", "null", "dielectric", "bumpmap"):
                self._bsdf_ids[name] = self._add_bsdf(v)
            elif t in ("homogeneous", "heterogeneous"):
                self._medium_ids[name] = self._add_medium(v, name)
            elif t in _TISSUE_MEDIA:
                self._medium_ids[name] = self._add_tissue_medium(v, name)
        for name, v in self._named.items():
            t = v.get("type")
            if t in ("rectangle", "cube", "obj", "sphere", "mesh"):
                self._add_shape(v, name)
            elif t in ("constant", "envmap"):
                self._add_env(v)
            elif t == "perspective":
                self._sensor_dicts = getattr(self, "_sensor_dicts", []) + [v]
            elif t in ("path", "volpath", "prbvolpath", "biovolpath", "volpathmis"):
                self._integrator_dict = v
            elif t in ("biovolpath06", "direct", "aov", "moment", "ptracer", "stokes", "depth",
                       "prb", "prbvolpath_basic", "direct_projective", "prb_projective"):
                raise NotImplementedError(f"integrator plugin '{t}' is outside the hot path (path, volpath, volpathmis, prbvolpath, biovolpath)")
            elif t in ("diffuse", "roughdielectric", "principled", "null", "dielectric", "bumpmap", "homogeneous", "heterogeneous") or t in _TISSUE_MEDIA:
                pass                                  # named BSDFs / media: handled by the first pass
            else:
                # anything else would silently vanish from the render (a `point` emitter, a `ply` shape, a `plastic` BSDF ...)
                raise NotImplementedError(f"plugin '{t}' (object '{name}') is outside the hot path; supported top-level plugins: "
                                          "perspective, path/volpath/volpathmis/prbvolpath/biovolpath, rectangle/cube/sphere/obj/mesh, constant/envmap, "
                                          "diffuse/roughdielectric/principled/null/dielectric/bumpmap, homogeneous/heterogeneous, "
                                          "liver/glissonCapsule/parenchyma")
        if not hasattr(self, "_sensor_dicts"):
            raise ValueError("scene has no 'perspective' sensor")
        # Scene::sensors(): every sensor of the scene, in order; render(scene, sensor=i) picks one (src/render/integrator.cpp:33-44)
        self.sensors, self._sensor_spp = [], []
        for v in self._sensor_dicts:
            kind0, seed0 = getattr(self, "sampler_kind", None), getattr(self, "base_seed", None)
            self._build_sensor(v)
            if kind0 is not None and (kind0 != self.sampler_kind or seed0 != self.base_seed):
                raise NotImplementedError("the sensors of one scene must share the sampler plugin and its seed on this path")
            self.sensors.append(self.sensor); self._sensor_spp.append(self.spp)
        self.select_sensor(0)
        self._build_integrator(getattr(self, "_integrator_dict", {"type": "path"}))
        self._ctx = None

    def select_sensor(self, index: int):
        """Make sensor `index` the one film_size / spp / pod() describe (the backend is told by Context.render(sensor=...))."""
        if not 0 <= index < len(self.sensors):
            raise RuntimeError(f"Scene::render(): sensor index {index} is out of bounds!")
        self.sensor_index = index; self.sensor = self.sensors[index]; self.spp = self._sensor_spp[index]

    # ---- BSDFs ------------------------------------------------------------
    def _resolve_bsdf(self, v):
        if v is None:
            return self._add_bsdf({"type": "diffuse"})        # Shape default BSDF (src/render/shape.cpp)
        if v.get("type") == "ref":
            return self._bsdf_ids[v["id"]]
        return self._add_bsdf(v)

    def _add_texture(self, v) -> int:
        """`bitmap` texture (src/textures/bitmap.cpp:172-290) -> texture slot (1 + index)."""
        if v.get("type") != "bitmap":
            raise NotImplementedError(f"texture plugin '{v.get('type')}' is outside the hot path")
        raw = bool(v.get("raw", False))
        if "filename" in v:
            px = read_texture(os.path.join(self.base_dir, v["filename"]), raw)
        elif "data" in v:                                      # tensor input is used as is (:236-262)
            px = np.ascontiguousarray(np.asarray(v["data"], dtype=np.float32))
            if px.ndim != 3:
                raise RuntimeError(f"Bitmap raw tensor has dimension {px.ndim}, expected 3")
        elif "bitmap" in v:
            px = texture_from_pixels(np.asarray(v["bitmap"]), raw)
        else:
            raise RuntimeError('bitmap: one of "filename", "data" or "bitmap" must be given')
        if px.shape[2] not in (1, 3):
            raise RuntimeError("bitmap: only textures with 1 or 3 channels are supported!")
        if px.shape[0] < 2 or px.shape[1] < 2:
            raise RuntimeError("bitmap: image must be at least 2x2 pixels in size")
        t = K.lr_texture(); self._keep.append(px)
        t.data = px.ctypes.data_as(K.c_float_p); t.height, t.width, t.channels = px.shape
        ft = str(v.get("filter_type", "bilinear")).lower()
        if ft not in ("bilinear", "nearest"):
            raise RuntimeError(f'Invalid filter type "{ft}", must be one of: "nearest", or "bilinear"!')
        t.filter = K.TEX_BILINEAR if ft == "bilinear" else K.TEX_NEAREST
        wm = str(v.get("wrap_mode", "repeat")).lower()
        if wm not in ("repeat", "mirror", "clamp"):
            raise RuntimeError(f'Invalid wrap mode "{wm}", must be one of: "repeat", "mirror", or "clamp"!')
        t.wrap = {"repeat": K.WRAP_REPEAT, "mirror": K.WRAP_MIRROR, "clamp": K.WRAP_CLAMP}[wm]
        t.to_uv = _uv6(v)
        self.textures.append(t)
        return len(self.textures)

    def _set_reflectance(self, b, r, default):
        """The colour texture slot of a BSDF block (diffuse `reflectance`, principled `base_color`): constant,
        `checkerboard` (src/textures/checkerboard.cpp) or `bitmap` (src/textures/bitmap.cpp)."""
        if isinstance(r, dict) and r.get("type") == "checkerboard":
            b.tex_kind = 1
            b.reflectance = _f3(_rgb(r.get("color0"), 0.4)); b.tex_color1 = _f3(_rgb(r.get("color1"), 0.2))
            m = _xf(r, "to_uv").matrix
            b.tex_to_uv = (C.c_float * 6)(m[0, 0], m[0, 1], m[0, 3], m[1, 0], m[1, 1], m[1, 3])
        elif isinstance(r, dict) and r.get("type") == "bitmap":
            b.tex_kind = 2; b.refl_tex = self._add_texture(r)
        else:
            b.reflectance = _f3(_rgb(r, default))

    def _add_bsdf(self, v):
        if v["type"] == "bumpmap":                             # src/bsdfs/bumpmap.cpp:101-127
            nested = [x for x in v.values() if isinstance(x, dict) and x.get("type") in
                      ("diffuse", "roughdielectric", "principled", "null", "dielectric", "ref", "bumpmap")]
            texs = [x for x in v.values() if isinstance(x, dict) and x.get("type") in ("bitmap", "checkerboard")]
            if len(nested) != 1:
                raise RuntimeError("Exactly one BSDF child object must be specified." if not nested
                                   else "Only a single BSDF child object can be specified.")
            if len(texs) != 1:
                raise RuntimeError("Exactly one Texture child object must be specified." if not texs
                                   else "Only a single Texture child object can be specified.")
            if nested[0]["type"] == "bumpmap":
                raise NotImplementedError("nested bumpmap BSDFs are outside the hot path")
            src = self.bsdfs[self._resolve_bsdf(nested[0])]
            b = K.lr_bsdf.from_buffer_copy(src)
            b.bump_tex = self._add_texture(texs[0]); b.bump_scale = _flt(v.get("scale"), 1.0)
            b.bump_flags = (K.BUMP_FLIP_INVALID_NORMALS if v.get("flip_invalid_normals", True) else 0) | \
                           (K.BUMP_SHADOWING_FUNCTION if v.get("use_shadowing_function", True) else 0)
            self.bsdfs.append(b)
            return len(self.bsdfs) - 1
        b = K.lr_bsdf(); t = v["type"]
        b.sample_visible = 1; b.eta = 1.0
        b.reflectance = _f3([1, 1, 1]); b.transmittance = _f3([1, 1, 1])
        if t == "diffuse":                                     # diffuse.cpp: reflectance default 0.5
            b.kind = K.BSDF_DIFFUSE
            self._set_reflectance(b, v.get("reflectance"), 0.5)
        elif t == "null":
            b.kind = K.BSDF_NULL
        elif t in ("dielectric", "roughdielectric"):           # roughdielectric.cpp:165-220
            b.kind = K.BSDF_DIELECTRIC if t == "dielectric" else K.BSDF_ROUGHDIELECTRIC
            int_ior = _ior(v.get("int_ior"), "bk7"); ext_ior = _ior(v.get("ext_ior"), "air")
            if int_ior < 0 or ext_ior < 0 or int_ior == ext_ior:
                raise RuntimeError("The interior and exterior indices of refraction must be positive and differ!")
            b.eta = int_ior / ext_ior
            b.reflectance = _f3(_rgb(v.get("specular_reflectance"), 1.0))
            b.transmittance = _f3(_rgb(v.get("specular_transmittance"), 1.0))
            if t == "roughdielectric":
                distr = str(v.get("distribution", "beckmann")).lower()
                if distr not in ("beckmann", "ggx"):
                    raise RuntimeError(f'Specified an invalid distribution "{distr}", must be "beckmann" or "ggx"!')
                b.mf_kind = K.MF_BECKMANN if distr == "beckmann" else K.MF_GGX
                b.sample_visible = 1 if v.get("sample_visible", True) else 0
                if "alpha_u" in v or "alpha_v" in v:
                    b.alpha_u = _flt(v["alpha_u"]); b.alpha_v = _flt(v["alpha_v"])
                else:
                    b.alpha_u = b.alpha_v = _flt(v.get("alpha"), 0.1)
        elif t == "principled":                                # principled.cpp:190-238
            b.kind = K.BSDF_PRINCIPLED
            self._set_reflectance(b, v.get("base_color"), 0.5); b.roughness = _flt(v.get("roughness"), 0.5)
            flags = 0
            def opt(name, bit, default=0.0):
                nonlocal flags
                val = _flt(v.get(name), default)
                if name in v and not (isinstance(v[name], (int, float)) and float(v[name]) == 0.0):   # get_flag()
                    flags |= bit
                return val
            b.anisotropic = opt("anisotropic", K.PR_HAS_ANISOTROPIC); b.spec_trans = opt("spec_trans", K.PR_HAS_SPEC_TRANS)
            b.sheen = opt("sheen", K.PR_HAS_SHEEN); b.sheen_tint = opt("sheen_tint", K.PR_HAS_SHEEN_TINT)
            b.flatness = opt("flatness", K.PR_HAS_FLATNESS); b.spec_tint = opt("spec_tint", K.PR_HAS_SPEC_TINT)
            b.metallic = opt("metallic", K.PR_HAS_METALLIC); b.clearcoat = opt("clearcoat", K.PR_HAS_CLEARCOAT)
            b.clearcoat_gloss = _flt(v.get("clearcoat_gloss"), 0.0)
            b.main_spec_srate = _flt(v.get("main_specular_sampling_rate"), 1.0)
            b.clearcoat_srate = _flt(v.get("clearcoat_sampling_rate"), 1.0)
            b.diff_refl_srate = _flt(v.get("diffuse_reflectance_sampling_rate"), 1.0)
            has_st = bool(flags & K.PR_HAS_SPEC_TRANS)
            if "eta" in v and "specular" in v:
                raise RuntimeError('Specified an invalid index of refraction property, either use "eta" or "specular" !')
            if "eta" in v:
                flags |= K.PR_ETA_SPECULAR; eta = np.float32(v["eta"])
                if has_st and eta == 1:
                    eta = np.float32(1.001)
                b.eta = float(eta)
            else:
                spec = np.float32(v.get("specular", 0.5))
                if has_st and spec == 0:
                    spec = np.float32(1e-3)
                b.specular = float(spec)
                b.eta = float(np.float32(2.0) * (np.float32(1.0) / (np.float32(1.0) - np.sqrt(np.float32(0.08) * spec))) - np.float32(1.0))
            b.flags = flags
        else:
            raise NotImplementedError(f"BSDF plugin '{t}' is outside the hot path")
        self.bsdfs.append(b)
        return len(self.bsdfs) - 1

    # ---- volumes / media ----------------------------------------------------
    def _add_grid(self, v, differentiable=False):
        """gridvolume (src/volumes/grid.cpp:164-307): data [z][y][x][c]."""
        if "filename" in v:
            data = read_vol(os.path.join(self.base_dir, v["filename"]))
        else:
            data = np.asarray(v.get("grid", v.get("data")), dtype=np.float32)
        if data.ndim == 3:
            data = data[..., None]
        data = np.ascontiguousarray(data, dtype=np.float32)
        g = K.lr_grid(); rz, ry, rx, ch = data.shape
        if ch not in (1, 3):
            raise RuntimeError("gridvolume: only 1 or 3 channels are supported on this path")
        g.res = (C.c_uint32 * 3)(rx, ry, rz); g.channels = ch
        to_world = _xf(v); to_local = to_world.inverse()
        g.to_local = _f16(to_local.matrix)
        corners = np.array([[x, y, z] for x in (0, 1) for y in (0, 1) for z in (0, 1)], dtype=np.float64)
        wc = to_world.transform_affine(corners).astype(np.float32)
        g.bbox_min = _f3(wc.min(0)); g.bbox_max = _f3(wc.max(0))
        ft = str(v.get("filter_type", "trilinear")).lower(); wm = str(v.get("wrap_mode", "clamp")).lower()
        g.filter = {"nearest": K.FILTER_NEAREST, "trilinear": K.FILTER_TRILINEAR}[ft]
        g.wrap = {"repeat": K.WRAP_REPEAT, "mirror": K.WRAP_MIRROR, "clamp": K.WRAP_CLAMP}[wm]
        g.max_value = float(v["max_value"]) if "max_value" in v else -1.0
        g.differentiable = 1 if differentiable else 0
        g.data = data.ctypes.data_as(K.c_float_p)
        self._keep.append(data); self.grids.append(g); self.grid_arrays.append(data)
        return len(self.grids) - 1

    def _add_medium(self, v, name="medium"):
        """homogeneous.cpp:137-181 / heterogeneous.cpp:155-201 / medium.cpp:16-32."""
        m = K.lr_medium(); t = v["type"]
        m.kind = K.MEDIUM_HOMOGENEOUS if t == "homogeneous" else K.MEDIUM_HETEROGENEOUS
        m.scale = _flt(v.get("scale"), 1.0)
        m.sample_emitters = 1 if v.get("sample_emitters", True) else 0
        m.has_spectral_extinction = 1 if v.get("has_spectral_extinction", True) else 0
        ph = v.get("phase", {"type": "isotropic"})
        if ph["type"] == "isotropic":
            m.phase = K.PHASE_ISOTROPIC
        elif ph["type"] == "hg":
            m.phase = K.PHASE_HG; m.phase_g = _flt(ph.get("g"), 0.8)
            if not (-1 < m.phase_g < 1):
                raise RuntimeError("The asymmetry parameter must lie in the interval (-1, 1)!")
        else:
            raise NotImplementedError(f"phase function '{ph['type']}' is outside the hot path")
        m.sigma_t_grid = -1; m.albedo_grid = -1
        m.sigma_t_const = _f3([1, 1, 1]); m.albedo_const = _f3([0.75] * 3)
        for key, dst_grid, dst_const, default in (("sigma_t", "sigma_t_grid", "sigma_t_const", 1.0),
                                                  ("albedo", "albedo_grid", "albedo_const", 0.75)):
            val = v.get(key, default)
            if isinstance(val, dict) and val.get("type") == "gridvolume":
                if t == "homogeneous":
                    raise RuntimeError("homogeneous medium: grid volumes are not supported, use 'heterogeneous'")
                # both grids are differentiable parameters of the medium (heterogeneous.cpp:177-186: "albedo", "sigma_t")
                gi = self._add_grid(val, differentiable=(key == "sigma_t" or self._integrator_type == "prbvolpath"))
                setattr(m, dst_grid, gi)
                self.param_keys[f"{name}.{key}.data"] = gi
            elif isinstance(val, dict) and val.get("type") == "constvolume":
                setattr(m, dst_const, _f3(_rgb(val.get("value"), default)))
            else:
                setattr(m, dst_const, _f3(_rgb(val, default)))
        if t == "heterogeneous" and m.sigma_t_grid < 0:
            # constant sigma_t inside a heterogeneous medium: 1-voxel... represent with a 2^3 grid
            c = _rgb(v.get("sigma_t", 1.0))
            gi = self._add_grid({"grid": np.full((2, 2, 2, 3), 1.0, dtype=np.float32) * c})
            m.sigma_t_grid = gi
        self.media.append(m)
        return len(self.media) - 1

    def _add_tissue_medium(self, v, name="medium"):
        """The fork's tissue media: liver.cpp:139-201, glissonCapsule.cpp:138-186, parenchyma.cpp:139-152.  The POD carries the
        values as the constructors store them, including their property-name mix-up: G is read from "..._B" and B from
        "..._G" for every collagen layer and for elastin layers 1 and 2 (liver.cpp:149-150 ... 174-175)."""
        m = K.lr_medium(); t = v["type"]
        m.kind = _TISSUE_MEDIA[t]
        m.scale = _flt(v.get("scale"), 1.0)
        m.sample_emitters = 1 if v.get("sample_emitters", t != "parenchyma") else 0          # unused by biovolpath
        m.has_spectral_extinction = 1 if v.get("has_spectral_extinction", t != "parenchyma") else 0   # parenchyma.cpp:150
        ph = v.get("phase", {"type": "isotropic"})
        if ph["type"] == "isotropic":
            m.phase = K.PHASE_ISOTROPIC
        elif ph["type"] == "hg":
            m.phase = K.PHASE_HG; m.phase_g = _flt(ph.get("g"), 0.8)
        else:
            raise NotImplementedError(f"phase function '{ph['type']}' is outside the hot path")
        m.sigma_t_grid = -1; m.albedo_grid = -1
        for key, default in (("sigma_t", 1.0), ("albedo", 0.75)):
            val = v.get(key, default)
            if isinstance(val, dict):
                if val.get("type") != "constvolume":
                    raise NotImplementedError(f"{t} medium: '{key}' must be constant (the fork's scenes use none)")
                val = val.get("value")
            setattr(m, key + "_const", _f3(_rgb(val, default)))
        lim = [_flt(v.get(f"layer{i}Limit"), d) for i, d in zip((1, 2, 3, 4), (0.0065, 0.0072, 0.0083, 0.01))]
        m.layer_limit = (C.c_float * 4)(*lim)
        for L in range(4):
            for arr, nm, swapped in ((m.sigma_collagen, "collagen", True), (m.sigma_elastin, "elastin", L < 2)):
                r = _flt(v.get(f"sigma_{nm}{L + 1}_R"), 1.0); g = _flt(v.get(f"sigma_{nm}{L + 1}_G"), 1.0); b = _flt(v.get(f"sigma_{nm}{L + 1}_B"), 1.0)
                arr[L][0], arr[L][1], arr[L][2] = (r, b, g) if swapped else (r, g, b)
        for key in ("sigma_blood", "sigma_bile", "sigma_lipid_water"):
            val = v.get(key, 1.0)
            if isinstance(val, dict):
                if val.get("type") not in ("constvolume", "rgb"):
                    raise NotImplementedError(f"{t} medium: '{key}' must be constant")
                val = val.get("value")
            setattr(m, key, _f3(_rgb(val, 1.0)))
        m.sigma_hepatocity = _flt(v.get("sigma_hepatocity"), 1.0)
        self.media.append(m)
        return len(self.media) - 1

    def _resolve_medium(self, v, name):
        if v is None:
            return -1
        if v.get("type") == "ref":
            return self._medium_ids[v["id"]]
        if v.get("type") in _TISSUE_MEDIA:
            return self._add_tissue_medium(v, name)
        return self._add_medium(v, name)

    # ---- shapes ---------------------------------------------------------------
    def _add_shape(self, v, name):
        t = v["type"]; to_world = _xf(v)
        if t == "rectangle":
            md = S.rectangle(to_world)
        elif t == "cube":
            md = S.cube(to_world)
        elif t == "sphere":
            md = S.icosphere(to_world, int(v.get("subdivisions", 5)), v.get("center", (0, 0, 0)), float(v.get("radius", 1.0)))
        elif t == "obj":
            md = S.load_obj(os.path.join(self.base_dir, v["filename"]), to_world, bool(v.get("face_normals", False)))
        else:  # 'mesh': raw arrays
            md = S.MeshData(to_world.transform_affine(np.asarray(v["positions"], dtype=np.float64)), v["indices"],
                            None if v.get("normals") is None else S._xf_normals(to_world, v["normals"]), v.get("texcoords"))
        m = K.lr_mesh()
        m.positions = md.positions.ctypes.data_as(K.c_float_p); m.indices = md.indices.ctypes.data_as(K.c_uint32_p)
        m.normals = md.normals.ctypes.data_as(K.c_float_p) if md.normals is not None else None
        m.texcoords = md.texcoords.ctypes.data_as(K.c_float_p) if md.texcoords is not None else None
        m.n_vertices = md.positions.shape[0]; m.n_faces = md.indices.shape[0]
        m.bsdf = self._resolve_bsdf(v.get("bsdf"))
        m.interior_medium = self._resolve_medium(v.get("interior"), f"{name}.interior_medium")
        m.exterior_medium = self._resolve_medium(v.get("exterior"), f"{name}.exterior_medium")
        m.emitter = -1; m.flip_normals = 0
        self._keep.append(md)
        mesh_id = len(self.meshes)
        em = v.get("emitter")
        if em is not None:
            if em["type"] != "area":
                raise NotImplementedError("only 'area' emitters can be attached to shapes")
            e = K.lr_emitter(); e.kind = K.EMITTER_AREA; e.mesh = mesh_id
            e.radiance = _f3(_rgb(em.get("radiance"), 1.0)); e.scale = 1.0
            if t == "rectangle":
                o, eu, ev, n, inv_area = S.rectangle_emitter_params(to_world)
                e.is_rectangle = 1; e.rect_origin = _f3(o); e.rect_edge_u = _f3(eu); e.rect_edge_v = _f3(ev)
                e.rect_normal = _f3(n); e.inv_area = inv_area
            m.emitter = len(self.emitters); self.emitters.append(e)
        self.meshes.append(m)

    def _add_env(self, v):
        e = K.lr_emitter(); e.mesh = -1; e.scale = 1.0
        if v["type"] == "constant":
            e.kind = K.EMITTER_CONSTANT; e.radiance = _f3(_rgb(v.get("radiance"), 1.0))
        else:                                                   # envmap.cpp:113-232
            e.kind = K.EMITTER_ENVMAP
            if "filename" in v:
                from .image_io import read_image_rgb
                px = read_image_rgb(os.path.join(self.base_dir, v["filename"]))
            else:
                px = np.asarray(v["bitmap"], dtype=np.float32)
            px = np.ascontiguousarray(px[..., :3], dtype=np.float32)
            e.height, e.width = px.shape[0], px.shape[1]
            e.pixels = px.ctypes.data_as(K.c_float_p); self._keep.append(px)
            e.scale = _flt(v.get("scale"), 1.0)
            e.to_world = _f16(_xf(v).matrix)
        self.emitters.append(e)

    # ---- sensor / film / sampler / integrator ---------------------------------
    def _build_sensor(self, v):
        """perspective.cpp + hdrfilm.cpp + independent.cpp property names."""
        s = K.lr_sensor(); film = v.get("film", {})
        W = int(film.get("width", 768)); H = int(film.get("height", 576))
        to_world = _xf(v)
        if to_world.has_scale():
            raise RuntimeError("Scale factors in the camera-to-world transformation are not allowed!")
        near = float(v.get("near_clip", 1e-2)); far = float(v.get("far_clip", 1e4))
        if "fov" in v:
            fov_x = parse_fov(float(v["fov"]), str(v.get("fov_axis", "x")), W / H)
        else:                                                   # focal_length default "50mm" (src/render/sensor.cpp:160-174)
            fl = float(str(v.get("focal_length", "50mm")).replace("mm", ""))
            import math
            fov_x = parse_fov(2.0 * math.degrees(math.atan(math.sqrt(36.0 * 36 + 24 * 24) / (2.0 * fl))), "diagonal", W / H)
        # crop window (Film::Film, src/render/film.cpp:30-52): crop_offset_x/_y, crop_width/_height; the render covers the crop only
        cw = int(film.get("crop_width", W)); chh = int(film.get("crop_height", H))
        cx = int(film.get("crop_offset_x", 0)); cy = int(film.get("crop_offset_y", 0))
        if cx < 0 or cy < 0 or cw <= 0 or chh <= 0 or cx + cw > W or cy + chh > H:
            raise RuntimeError("Invalid crop window specification!")
        c2s = perspective_projection((W, H), (cw, chh), (cx, cy), fov_x, near, far)
        s.to_world = _f16(to_world.matrix); s.sample_to_camera = _f16(c2s.inverse().matrix)
        s.near_clip = near; s.far_clip = far; s.film_width = cw; s.film_height = chh; s.crop_offset_x = cx; s.crop_offset_y = cy
        rf = film.get("rfilter", {"type": "gaussian"}); rt = rf.get("type", "gaussian")
        if rt == "box":
            s.rfilter = K.RFILTER_BOX; s.rfilter_radius = 0.5
        elif rt == "tent":
            s.rfilter = K.RFILTER_TENT; s.rfilter_radius = float(rf.get("radius", 1.0))
        elif rt == "gaussian":
            s.rfilter = K.RFILTER_GAUSSIAN; s.rfilter_stddev = float(rf.get("stddev", 0.5)); s.rfilter_radius = 4 * s.rfilter_stddev
        else:
            raise NotImplementedError(f"rfilter '{rt}' is outside the hot path")
        pf = str(film.get("pixel_format", "rgb")).lower()
        if pf not in ("rgb", "rgba"):
            raise NotImplementedError("hdrfilm pixel_format must be rgb or rgba on this path")
        s.has_alpha = 1 if pf == "rgba" else 0
        s.medium = self._resolve_medium(v.get("medium"), "sensor.medium")
        smp = v.get("sampler", {"type": "independent"})
        st = smp.get("type", "independent")
        if st not in ("independent", "ldsampler"):
            raise NotImplementedError("only the 'independent' and 'ldsampler' samplers are on the hot path")
        self.sampler_kind = K.SAMPLER_LD if st == "ldsampler" else K.SAMPLER_INDEPENDENT
        self.spp = int(smp.get("sample_count", 4)); self.base_seed = int(smp.get("seed", 0))
        if self.sampler_kind == K.SAMPLER_LD:
            self.spp = ld_sample_count(self.spp)
        self.sensor = s

    def _build_integrator(self, v):
        i = K.lr_integrator()
        i.kind = {"path": K.INTEGRATOR_PATH, "volpath": K.INTEGRATOR_VOLPATH, "prbvolpath": K.INTEGRATOR_PRBVOLPATH,
                  "biovolpath": K.INTEGRATOR_BIOVOLPATH, "volpathmis": K.INTEGRATOR_VOLPATHMIS}[v["type"]]
        i.spectral_mis = 1 if v.get("use_spectral_mis", True) else 0                          # volpathmis.cpp:79
        md = int(v.get("max_depth", 6 if v["type"] == "prbvolpath" else -1))
        if md < 0 and md != -1:
            raise RuntimeError('"max_depth" must be set to -1 (infinite) or a value >= 0')
        i.max_depth = md; i.rr_depth = int(v.get("rr_depth", 5))
        if i.rr_depth <= 0:
            raise RuntimeError('"rr_depth" must be set to a value greater than zero!')
        i.hide_emitters = 1 if v.get("hide_emitters", False) else 0
        i.tracking = {"reference": K.TRACKING_REFERENCE, "dda": K.TRACKING_DDA}[str(v.get("tracking", "reference")).lower()]
        i.base_seed = self.base_seed; i.majorant_cell = int(v.get("majorant_cell", 0)); i.sampler = getattr(self, "sampler_kind", K.SAMPLER_INDEPENDENT)
        self.integrator = i; self.integrator_name = v["type"]

    # ---- POD ---------------------------------------------------------------------
    def pod(self) -> K.lr_scene:
        sc = K.lr_scene(); sc.api_version = K.LR_API_VERSION
        def arr(T, lst):
            a = (T * max(len(lst), 1))(*lst); self._keep.append(a); return a
        sc.n_meshes, sc.n_bsdfs, sc.n_media = len(self.meshes), len(self.bsdfs), len(self.media)
        sc.n_grids, sc.n_emitters, sc.n_textures = len(self.grids), len(self.emitters), len(self.textures)
        sc.textures = arr(K.lr_texture, self.textures)
        sc.meshes = arr(K.lr_mesh, self.meshes); sc.bsdfs = arr(K.lr_bsdf, self.bsdfs); sc.media = arr(K.lr_medium, self.media)
        sc.grids = arr(K.lr_grid, self.grids); sc.emitters = arr(K.lr_emitter, self.emitters)
        sc.sensor = self.sensor; sc.integrator = self.integrator
        return sc

    @property
    def film_size(self):
        return int(self.sensor.film_width), int(self.sensor.film_height)

    @property
    def out_channels(self):
        return 4 if self.sensor.has_alpha else 3


def load_dict(d: dict, base_dir: str = ".") -> Scene:
    """`mi.load_dict` for the hot-path plugin subset."""
    return Scene(d, base_dir)
