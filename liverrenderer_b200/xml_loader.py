"""`mi.load_file`: the Mitsuba 3 XML scene format (src/core/xml.cpp) for the plugin subset of the hot path, turned into
the dictionary `load_dict` takes.  What the fork's scene files use is covered: `<default>` parameters with `$name`
substitution (overridable through keyword arguments, as `mi.load_file(path, spp=…)`), `<include>`, `<ref>`, nested
plugin objects, scalar / colour / vector properties and `<transform>` blocks whose operations apply in file order
(each one multiplies from the left, xml.cpp `Transform` handling)."""
from __future__ import annotations

import os
import re
import xml.etree.ElementTree as ET
import numpy as np

from .transforms import ScalarTransform4f
from .scene import Scene

_OBJECT_TAGS = {"scene", "integrator", "sensor", "film", "sampler", "rfilter", "shape", "bsdf", "medium", "phase",
                "emitter", "texture", "volume", "spectrum_object"}
# key a nested object gets when the file gives it no `name` (the reference matches such children by class)
_DEFAULT_KEY = {"film": "film", "sampler": "sampler", "rfilter": "rfilter", "bsdf": "bsdf", "phase": "phase",
                "emitter": "emitter", "integrator": "integrator", "sensor": "sensor"}


def _floats(s: str):
    return [float(x) for x in re.split(r"[\s,]+", s.strip()) if x]


class _Loader:
    def __init__(self, params, root_dir):
        self.params = dict(params); self.counter = 0
        self.paths = [root_dir]                 # FileResolver: the scene's directory, then <path> entries and include directories

    def resolve(self, name: str) -> str:
        if os.path.isabs(name):
            return name
        for p in self.paths:
            c = os.path.join(p, name)
            if os.path.exists(c):
                return os.path.abspath(c)
        return name

    def subst(self, s: str) -> str:
        if "$" not in s:
            return s
        def rep(m):
            k = m.group(1)
            if k not in self.params:
                raise RuntimeError(f'undefined parameter "${k}" (add <default name="{k}" …/> or pass {k}=… to load_file)')
            return str(self.params[k])
        return re.sub(r"\$(\w+)", rep, s)

    def attr(self, e, k, default=None):
        v = e.get(k)
        return default if v is None else self.subst(v)

    def vec3(self, e, default=0.0):
        if e.get("value") is not None:
            v = _floats(self.attr(e, "value"))
            return v * 3 if len(v) == 1 else v
        return [float(self.attr(e, a, default)) for a in ("x", "y", "z")]

    def transform(self, e) -> ScalarTransform4f:
        m = ScalarTransform4f()
        for op in e:
            t = op.tag
            if t == "translate":
                o = ScalarTransform4f().translate(self.vec3(op, 0.0))
            elif t == "scale":
                o = ScalarTransform4f().scale(self.vec3(op, 1.0))
            elif t == "rotate":
                axis = [float(self.attr(op, a, 0.0)) for a in ("x", "y", "z")] if op.get("value") is None else _floats(self.attr(op, "value"))
                o = ScalarTransform4f().rotate(axis, float(self.attr(op, "angle")))
            elif t == "lookat":
                o = ScalarTransform4f().look_at(_floats(self.attr(op, "origin")), _floats(self.attr(op, "target")),
                                                _floats(self.attr(op, "up", "0, 1, 0")))
            elif t == "matrix":
                v = _floats(self.attr(op, "value"))
                if len(v) == 9:
                    mm = np.eye(4); mm[:3, :3] = np.reshape(v, (3, 3)); v = mm.reshape(-1)
                o = ScalarTransform4f(np.reshape(v, (4, 4)))
            else:
                raise RuntimeError(f'unknown transform operation "<{t}>"')
            m = o @ m                                           # later operations apply after earlier ones
        return m

    def prop(self, e, base_dir):
        t = e.tag
        if t == "float":
            return float(self.attr(e, "value"))
        if t == "integer":
            return int(self.attr(e, "value"))
        if t == "boolean":
            return self.attr(e, "value").strip().lower() == "true"
        if t == "string":
            v = self.attr(e, "value")
            return self.resolve(v) if e.get("name") == "filename" else v
        if t == "rgb":
            return {"type": "rgb", "value": (_floats(self.attr(e, "value")) * 3)[:3] if len(_floats(self.attr(e, "value"))) == 1 else _floats(self.attr(e, "value"))}
        if t == "spectrum":
            v = _floats(self.attr(e, "value"))
            if len(v) != 1:
                raise NotImplementedError("only uniform <spectrum value=\"x\"/> entries are on the hot path (rgb variants)")
            return {"type": "rgb", "value": [v[0]] * 3}
        if t in ("vector", "point"):
            return self.vec3(e)
        if t == "transform":
            return self.transform(e)
        if t == "ref":
            return {"type": "ref", "id": self.attr(e, "id")}
        raise RuntimeError(f'unsupported tag "<{t}>"')

    def children(self, e, d, base_dir, top_level):
        for c in e:
            t = c.tag
            if t == "default":
                self.params.setdefault(c.get("name"), self.subst(c.get("value")))
            elif t == "alias":
                continue
            elif t == "path":                                     # <path value="…"/>: one more search directory
                self.paths.append(os.path.normpath(os.path.join(base_dir, self.attr(c, "value"))))
            elif t == "include":
                p = self.resolve(self.attr(c, "filename"))
                if not os.path.isabs(p):
                    p = os.path.join(base_dir, p)
                root = ET.parse(p).getroot()
                if os.path.dirname(p) not in self.paths:
                    self.paths.append(os.path.dirname(p))
                self.children(root, d, os.path.dirname(p), top_level)
            elif t in _OBJECT_TAGS:
                key = c.get("name") or (c.get("id") if top_level else None) or _DEFAULT_KEY.get(t)
                if key is None or key in d:
                    self.counter += 1; key = f"{key or t}_{self.counter}"
                d[key] = self.obj(c, base_dir)
            else:
                name = c.get("name")
                if name is None and t == "ref":                  # un-named reference: matched by the class of its target
                    self.counter += 1; name = f"ref_{self.counter}"
                d[name] = self.prop(c, base_dir)

    def obj(self, e, base_dir, top_level=False):
        d = {"type": self.attr(e, "type", "scene") if e.tag != "scene" else "scene"}
        self.children(e, d, base_dir, top_level)
        if e.tag == "integrator" and d["type"] in ("moment", "aov"):      # AOV wrappers around the radiance integrator
            nested = [v for v in d.values() if isinstance(v, dict) and v.get("type") not in (None, "ref")]
            if len(nested) == 1:
                return nested[0]
        return d


def xml_to_dict(path: str, **params) -> dict:
    """Parse a scene file into a `load_dict` dictionary.  Keyword arguments override `<default>` parameters."""
    root = ET.parse(path).getroot()
    if root.tag != "scene":
        raise RuntimeError(f'"{path}": root element must be <scene>')
    root_dir = os.path.dirname(os.path.abspath(path))
    ld = _Loader(params, root_dir)
    d = ld.obj(root, root_dir, top_level=True)
    # un-named references inside a shape: a BSDF / medium by the kind of the object they point to
    kinds = {k: v.get("type") for k, v in d.items() if isinstance(v, dict)}
    for v in d.values():
        if not isinstance(v, dict):
            continue
        for k in [k for k in v if k.startswith("ref_")]:
            target = kinds.get(v[k]["id"])
            if target in ("homogeneous", "heterogeneous"):
                v.setdefault("interior", v.pop(k))
            elif target is not None:
                v.setdefault("bsdf", v.pop(k))
    return d


def load_file(path: str, **params) -> Scene:
    return Scene(xml_to_dict(path, **params), os.path.dirname(os.path.abspath(path)))
