"""Host-side 4x4 transforms mirroring `mi.ScalarTransform4f`.

Reference: include/mitsuba/core/transform.h (translate/scale/rotate/look_at:177-205,
perspective:393-410) and include/mitsuba/render/sensor.h:234-269
(`perspective_projection`).  Everything is float64 on the host and cast to
float32 when flattened into the POD scene.
"""
from __future__ import annotations

import math
import numpy as np


class ScalarTransform4f:
    """Chainable affine/projective transform; `T().translate(a).rotate(..)` composes
    on the right exactly like the reference (`self @ other`)."""

    def __init__(self, matrix=None):
        self.matrix = np.eye(4) if matrix is None else np.array(matrix, dtype=np.float64).reshape(4, 4)

    # -- composition ---------------------------------------------------------
    def __matmul__(self, other):
        if isinstance(other, ScalarTransform4f):
            return ScalarTransform4f(self.matrix @ other.matrix)
        v = np.asarray(other, dtype=np.float64)
        return self.transform_affine(v)

    def inverse(self):
        return ScalarTransform4f(np.linalg.inv(self.matrix))

    def transform_affine(self, p):
        p = np.asarray(p, dtype=np.float64)
        r = p @ self.matrix[:3, :3].T + self.matrix[:3, 3]
        return r

    def transform_vector(self, v):
        return np.asarray(v, dtype=np.float64) @ self.matrix[:3, :3].T

    def transform_normal(self, n):
        return np.asarray(n, dtype=np.float64) @ np.linalg.inv(self.matrix[:3, :3])

    def translation(self):
        return self.matrix[:3, 3].copy()

    # -- builders (instance methods chain on the right, like the reference) --
    def translate(self, v):
        m = np.eye(4); m[:3, 3] = np.asarray(v, dtype=np.float64)
        return ScalarTransform4f(self.matrix @ m)

    def scale(self, v):
        v = np.asarray(v, dtype=np.float64)
        if v.ndim == 0:
            v = np.array([v, v, v])
        m = np.diag([v[0], v[1], v[2], 1.0])
        return ScalarTransform4f(self.matrix @ m)

    def rotate(self, axis, angle):
        """Rotation by `angle` degrees around `axis` (transform.h rotate())."""
        a = np.asarray(axis, dtype=np.float64); a = a / np.linalg.norm(a)
        s, c = math.sin(math.radians(angle)), math.cos(math.radians(angle))
        x, y, z = a
        m = np.eye(4)
        m[0, 0] = x * x + (1 - x * x) * c; m[0, 1] = x * y * (1 - c) - z * s; m[0, 2] = x * z * (1 - c) + y * s
        m[1, 0] = x * y * (1 - c) + z * s; m[1, 1] = y * y + (1 - y * y) * c; m[1, 2] = y * z * (1 - c) - x * s
        m[2, 0] = x * z * (1 - c) - y * s; m[2, 1] = y * z * (1 - c) + x * s; m[2, 2] = z * z + (1 - z * z) * c
        return ScalarTransform4f(self.matrix @ m)

    def look_at(self, origin, target, up):
        origin = np.asarray(origin, dtype=np.float64); target = np.asarray(target, dtype=np.float64)
        up = np.asarray(up, dtype=np.float64)
        d = target - origin; d /= np.linalg.norm(d)
        left = np.cross(up, d); left /= np.linalg.norm(left)
        new_up = np.cross(d, left)
        m = np.eye(4)
        m[:3, 0] = left; m[:3, 1] = new_up; m[:3, 2] = d; m[:3, 3] = origin
        return ScalarTransform4f(self.matrix @ m)

    @staticmethod
    def perspective(fov, near, far):
        recip = 1.0 / (far - near)
        cot = 1.0 / math.tan(math.radians(fov * 0.5))
        m = np.diag([cot, cot, far * recip, 0.0])
        m[2, 3] = -near * far * recip
        m[3, 2] = 1.0
        return ScalarTransform4f(m)

    def has_scale(self):
        m = self.matrix[:3, :3]
        return not np.allclose(m @ m.T, np.eye(3), atol=1e-3)

    def __repr__(self):
        return f"ScalarTransform4f({self.matrix.tolist()})"


T = ScalarTransform4f


def parse_fov(fov: float, fov_axis: str, aspect: float) -> float:
    """x-axis field of view in degrees (src/render/sensor.cpp:142-200)."""
    axis = fov_axis.lower()
    if axis == "smaller":
        axis = "y" if aspect > 1 else "x"
    elif axis == "larger":
        axis = "x" if aspect > 1 else "y"
    if axis == "x":
        return fov
    if axis == "y":
        return math.degrees(2.0 * math.atan(math.tan(0.5 * math.radians(fov)) * aspect))
    if axis == "diagonal":
        diagonal = 2.0 * math.tan(0.5 * math.radians(fov))
        width = diagonal / math.sqrt(1.0 + 1.0 / (aspect * aspect))
        return math.degrees(2.0 * math.atan(width * 0.5))
    raise ValueError(f"parse_fov(): invalid fov_axis '{fov_axis}'")


def perspective_projection(film_size, crop_size, crop_offset, fov_x, near, far) -> ScalarTransform4f:
    """camera space -> sample space (include/mitsuba/render/sensor.h:234-269)."""
    fw, fh = float(film_size[0]), float(film_size[1])
    rel_size = (crop_size[0] / fw, crop_size[1] / fh)
    rel_off = (crop_offset[0] / fw, crop_offset[1] / fh)
    aspect = fw / fh
    t = ScalarTransform4f()
    return (t.scale([1.0 / rel_size[0], 1.0 / rel_size[1], 1.0])
             .translate([-rel_off[0], -rel_off[1], 0.0])
             .scale([-0.5, -0.5 * aspect, 1.0])
             .translate([-1.0, -1.0 / aspect, 0.0])) @ ScalarTransform4f.perspective(fov_x, near, far)
