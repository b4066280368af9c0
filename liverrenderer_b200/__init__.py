"""liverrenderer_b200 -- B200-native backend for the LiverRenderer (Mitsuba 3 fork)
`path` / `volpath` / `prbvolpath` hot path.

The public names mirror the reference's Python entry points
(src/python/python/util.py: `mi.load_dict`, `mi.render`:394-525,
`mi.cornell_box`:567-703, `mi.traverse`:270-340) so code written against the
reference reads the same:

    import liverrenderer_b200 as mi
    scene = mi.load_dict(mi.cornell_box())
    img = mi.render(scene, spp=64)                    # numpy [H, W, 3]

Everything below `Integrator::render` runs in hand-written CUDA behind the C ABI
in include/lr_api.h; there is no CPU fallback.
"""
from __future__ import annotations

import numpy as np

from .transforms import ScalarTransform4f, T
from .scene import Scene, load_dict
from .backend import Context
from .volume_io import read_vol, write_vol
from .image_io import write_bitmap, read_texture
from . import _capi

__all__ = ["ScalarTransform4f", "T", "Scene", "load_dict", "load_file", "Context", "render", "render_backward",
           "cornell_box", "traverse", "sample_tea_32", "read_vol", "write_vol", "write_bitmap", "read_texture", "variant", "set_variant"]

_variant = "cuda_b200_rgb"


def variant():
    return _variant


def set_variant(*names):
    """Accepted for source compatibility; this backend has exactly one variant
    (fp32 RGB on CUDA sm_100a).  CPU variants are refused: no CPU path exists here."""
    for n in names:
        if n.startswith("cuda") and n.endswith("rgb"):
            return
    raise ImportError(f"Requested an unsupported variant {names}; liverrenderer_b200 only provides CUDA RGB")


def sample_tea_32(v0: int, v1: int, rounds: int = 4):
    """include/mitsuba/core/random.h:76-90 (host-side: used to derive seed_grad, util.py:503-505)."""
    v0 &= 0xffffffff; v1 &= 0xffffffff; s = 0
    for _ in range(rounds):
        s = (s + 0x9e3779b9) & 0xffffffff
        v0 = (v0 + ((((v1 << 4) & 0xffffffff) + 0xa341316c) ^ ((v1 + s) & 0xffffffff) ^ ((v1 >> 5) + 0xc8013ea4))) & 0xffffffff
        v1 = (v1 + ((((v0 << 4) & 0xffffffff) + 0xad90777d) ^ ((v0 + s) & 0xffffffff) ^ ((v0 >> 5) + 0x7e95761e))) & 0xffffffff
    return v0, v1


def _context_for(scene: Scene, device: int = 0) -> Context:
    ctx = getattr(scene, "_ctx", None)
    if ctx is None or ctx.device != device:
        ctx = Context(device).upload(scene)
        scene._ctx = ctx
    return ctx


def render(scene: Scene, params=None, sensor=0, integrator=None, seed: int = 0, seed_grad: int = 0,
           spp: int = 0, spp_grad: int = 0, device: int = 0, out=None):
    """`mi.render` (src/python/python/util.py:394-525), primal part."""
    if params is not None and hasattr(params, "update"):
        params.update()
    ctx = _context_for(scene, device)
    if not isinstance(sensor, int):
        raise NotImplementedError("mi.render: `sensor` is the index of one of the scene's sensors on this path")
    ctx.set_sensor(sensor)
    return ctx.render(seed=seed, spp=spp or scene.spp, out=out)


def render_backward(scene: Scene, grad_in, seed: int = 0, seed_grad: int = 0, spp: int = 0, spp_grad: int = 0,
                    device: int = 0):
    """Adjoint of `mi.render` (util.py `_RenderOp.backward` -> RBIntegrator.render_backward,
    ad/integrators/common.py:625-783).  Accumulates into the grid-gradient buffers; read them
    through `traverse(scene)` / `params.grad(key)`."""
    if seed_grad == 0:
        seed_grad = sample_tea_32(seed, 1)[0]                      # util.py:503-505
    ctx = _context_for(scene, device)
    ctx.render_backward(grad_in, seed=seed_grad, spp=spp_grad or spp or scene.spp)


class SceneParameters:
    """`mi.traverse` subset: the differentiable grid volumes of the scene's media
    (keys like `<shape>.interior_medium.sigma_t.data`, src/volumes/grid.cpp:309-327)."""

    def __init__(self, scene: Scene, device: int = 0):
        self.scene = scene; self.device = device
        self._dirty = set()

    def keys(self):
        return list(self.scene.param_keys.keys())

    def __contains__(self, k):
        return k in self.scene.param_keys

    def __getitem__(self, k):
        return self.scene.grid_arrays[self.scene.param_keys[k]]

    def __setitem__(self, k, v):
        gi = self.scene.param_keys[k]
        arr = self.scene.grid_arrays[gi]
        arr[...] = np.asarray(v, dtype=np.float32).reshape(arr.shape)
        self._dirty.add(k)

    def set_dirty(self, k):
        self._dirty.add(k)

    def update(self):
        """params.update(): re-upload and recompute the majorant (heterogeneous.cpp:173-175)."""
        if not self._dirty:
            return
        ctx = _context_for(self.scene, self.device)
        for k in self._dirty:
            gi = self.scene.param_keys[k]
            ctx.grid_update(gi, self.scene.grid_arrays[gi])
        self._dirty.clear()

    def grad(self, k):
        """`dr.grad(params[k])` after render_backward."""
        gi = self.scene.param_keys[k]
        ctx = _context_for(self.scene, self.device)
        return ctx.grid_grad_read(gi).reshape(self.scene.grid_arrays[gi].shape)

    def zero_grad(self):
        ctx = _context_for(self.scene, self.device)
        for gi in self.scene.param_keys.values():
            ctx.grid_grad_zero(gi)


def traverse(scene: Scene, device: int = 0) -> SceneParameters:
    return SceneParameters(scene, device)


def cornell_box():
    """`mi.cornell_box()` (src/python/python/util.py:567-703)."""
    return {
        "type": "scene",
        "integrator": {"type": "path", "max_depth": 8},
        "sensor": {
            "type": "perspective", "fov_axis": "smaller", "near_clip": 0.001, "far_clip": 100.0, "focus_distance": 1000,
            "fov": 39.3077, "to_world": T().look_at(origin=[0, 0, 3.90], target=[0, 0, 0], up=[0, 1, 0]),
            "sampler": {"type": "independent", "sample_count": 64},
            "film": {"type": "hdrfilm", "width": 256, "height": 256, "rfilter": {"type": "gaussian"},
                     "pixel_format": "rgb", "component_format": "float32"},
        },
        "white": {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.885809, 0.698859, 0.666422]}},
        "green": {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.105421, 0.37798, 0.076425]}},
        "red": {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.570068, 0.0430135, 0.0443706]}},
        "light": {"type": "rectangle",
                  "to_world": T().translate([0.0, 0.99, 0.01]).rotate([1, 0, 0], 90).scale([0.23, 0.19, 0.19]),
                  "bsdf": {"type": "ref", "id": "white"},
                  "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [18.387, 13.9873, 6.75357]}}},
        "floor": {"type": "rectangle", "to_world": T().translate([0.0, -1.0, 0.0]).rotate([1, 0, 0], -90),
                  "bsdf": {"type": "ref", "id": "white"}},
        "ceiling": {"type": "rectangle", "to_world": T().translate([0.0, 1.0, 0.0]).rotate([1, 0, 0], 90),
                    "bsdf": {"type": "ref", "id": "white"}},
        "back": {"type": "rectangle", "to_world": T().translate([0.0, 0.0, -1.0]), "bsdf": {"type": "ref", "id": "white"}},
        "green-wall": {"type": "rectangle", "to_world": T().translate([1.0, 0.0, 0.0]).rotate([0, 1, 0], -90),
                       "bsdf": {"type": "ref", "id": "green"}},
        "red-wall": {"type": "rectangle", "to_world": T().translate([-1.0, 0.0, 0.0]).rotate([0, 1, 0], 90),
                     "bsdf": {"type": "ref", "id": "red"}},
        "small-box": {"type": "cube", "to_world": T().translate([0.335, -0.7, 0.38]).rotate([0, 1, 0], -17).scale(0.3),
                      "bsdf": {"type": "ref", "id": "white"}},
        "large-box": {"type": "cube", "to_world": T().translate([-0.33, -0.4, -0.28]).rotate([0, 1, 0], 18.25).scale([0.3, 0.61, 0.3]),
                      "bsdf": {"type": "ref", "id": "white"}},
    }


def load_file(path: str, **kwargs) -> Scene:
    """`mi.load_file` for the XML subset the hot path needs (see xml_loader.py)."""
    from .xml_loader import load_file as _lf
    return _lf(path, **kwargs)
