"""Host-side triangle-mesh construction for the shape plugins on the hot path.

Reference: src/shapes/cube.cpp:96-160 (vertex tables), src/shapes/rectangle.cpp
(the reference's `rectangle` is analytic; here it is two triangles because the
backend traces triangles only -- SURVEY.md 8a5), src/shapes/obj.cpp,
src/render/mesh.cpp (recompute_vertex_normals).  `sphere` is tessellated
(icosphere), a documented deviation: the reference's sphere is analytic.
"""
from __future__ import annotations

import numpy as np

from .transforms import ScalarTransform4f


class MeshData:
    def __init__(self, positions, indices, normals=None, texcoords=None):
        self.positions = np.ascontiguousarray(positions, dtype=np.float32).reshape(-1, 3)
        self.indices = np.ascontiguousarray(indices, dtype=np.uint32).reshape(-1, 3)
        self.normals = None if normals is None else np.ascontiguousarray(normals, dtype=np.float32).reshape(-1, 3)
        self.texcoords = None if texcoords is None else np.ascontiguousarray(texcoords, dtype=np.float32).reshape(-1, 2)

    def bbox(self):
        return self.positions.min(0), self.positions.max(0)


def _xf_normals(to_world: ScalarTransform4f, n):
    n = to_world.transform_normal(np.asarray(n, dtype=np.float64))
    return n / np.linalg.norm(n, axis=-1, keepdims=True)


def rectangle(to_world: ScalarTransform4f) -> MeshData:
    p = np.array([[-1, -1, 0], [1, -1, 0], [1, 1, 0], [-1, 1, 0]], dtype=np.float64)
    uv = np.array([[0, 0], [1, 0], [1, 1], [0, 1]], dtype=np.float64)
    idx = np.array([[0, 1, 2], [0, 2, 3]], dtype=np.uint32)
    return MeshData(to_world.transform_affine(p), idx, None, uv)


def rectangle_emitter_params(to_world: ScalarTransform4f):
    """origin / edges / normal / 1/area for Rectangle::sample_position (rectangle.cpp:181-199)."""
    o = to_world.transform_affine(np.zeros(3))
    eu = to_world.transform_vector([1, 0, 0]); ev = to_world.transform_vector([0, 1, 0])
    n = _xf_normals(to_world, [0, 0, 1])
    area = np.linalg.norm(np.cross(eu, ev)) * 4.0
    return o, eu, ev, n, 1.0 / area


_CUBE_V = np.array([
    [1, -1, -1], [1, -1, 1], [-1, -1, 1], [-1, -1, -1], [1, 1, -1], [-1, 1, -1], [-1, 1, 1], [1, 1, 1],
    [1, -1, -1], [1, 1, -1], [1, 1, 1], [1, -1, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1], [-1, -1, 1],
    [-1, -1, 1], [-1, 1, 1], [-1, 1, -1], [-1, -1, -1], [1, 1, -1], [1, -1, -1], [-1, -1, -1], [-1, 1, -1]], dtype=np.float64)
_CUBE_N = np.repeat(np.array([[0, -1, 0], [0, 1, 0], [1, 0, 0], [0, 0, 1], [-1, 0, 0], [0, 0, -1]], dtype=np.float64), 4, axis=0)
_CUBE_UV = np.tile(np.array([[0, 1], [1, 1], [1, 0], [0, 0]], dtype=np.float64), (6, 1))
_CUBE_F = np.array([[0, 1, 2], [3, 0, 2], [4, 5, 6], [7, 4, 6], [8, 9, 10], [11, 8, 10], [12, 13, 14], [15, 12, 14],
                    [16, 17, 18], [19, 16, 18], [20, 21, 22], [23, 20, 22]], dtype=np.uint32)


def cube(to_world: ScalarTransform4f) -> MeshData:
    return MeshData(to_world.transform_affine(_CUBE_V), _CUBE_F, _xf_normals(to_world, _CUBE_N), _CUBE_UV)


def icosphere(to_world: ScalarTransform4f, subdivisions: int = 5, center=(0, 0, 0), radius=1.0) -> MeshData:
    t = (1.0 + 5 ** 0.5) / 2.0
    v = [[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
         [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]]
    v = [np.array(x, dtype=np.float64) / np.linalg.norm(x) for x in v]
    f = [[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6], [7, 1, 8],
         [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]]
    for _ in range(subdivisions):
        cache = {}
        def mid(a, b):
            key = (min(a, b), max(a, b))
            if key not in cache:
                m = v[a] + v[b]; m /= np.linalg.norm(m); v.append(m); cache[key] = len(v) - 1
            return cache[key]
        nf = []
        for a, b, c in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [[a, ab, ca], [b, bc, ab], [c, ca, bc], [ab, bc, ca]]
        f = nf
    n = np.array(v)
    p = n * radius + np.asarray(center, dtype=np.float64)
    return MeshData(to_world.transform_affine(p), np.array(f, dtype=np.uint32), _xf_normals(to_world, n), None)


def compute_vertex_normals(positions, indices):
    """Angle-weighted vertex normals (Mesh::recompute_vertex_normals, src/render/mesh.cpp;
    'Computing Vertex Normals from Polygonal Facets', Thuermer & Wuethrich 1998)."""
    p = positions.astype(np.float64); n = np.zeros_like(p)
    for k in range(3):
        i0, i1, i2 = indices[:, k], indices[:, (k + 1) % 3], indices[:, (k + 2) % 3]
        d0 = p[i1] - p[i0]; d1 = p[i2] - p[i0]
        fn = np.cross(p[indices[:, 1]] - p[indices[:, 0]], p[indices[:, 2]] - p[indices[:, 0]])
        ln = np.linalg.norm(fn, axis=1, keepdims=True); ln[ln == 0] = 1
        fn = fn / ln
        c = np.einsum("ij,ij->i", d0, d1) / np.maximum(np.linalg.norm(d0, axis=1) * np.linalg.norm(d1, axis=1), 1e-30)
        ang = np.arccos(np.clip(c, -1, 1))
        np.add.at(n, i0, fn * ang[:, None])
    ln = np.linalg.norm(n, axis=1, keepdims=True); ln[ln == 0] = 1
    return (n / ln).astype(np.float32)


def load_obj(path: str, to_world: ScalarTransform4f, face_normals: bool = False) -> MeshData:
    """Wavefront OBJ subset (v / vt / vn / f with polygons fan-triangulated); unique
    vertices per (v, vt, vn) key like src/shapes/obj.cpp."""
    vs, vts, vns, keys, tris = [], [], [], {}, []
    out_v, out_t, out_n = [], [], []
    with open(path, "r", errors="ignore") as fh:
        for line in fh:
            s = line.split()
            if not s:
                continue
            if s[0] == "v":
                vs.append([float(x) for x in s[1:4]])
            elif s[0] == "vt":
                vts.append([float(s[1]), float(s[2]) if len(s) > 2 else 0.0])
            elif s[0] == "vn":
                vns.append([float(x) for x in s[1:4]])
            elif s[0] == "f":
                ids = []
                for tok in s[1:]:
                    parts = tok.split("/")
                    vi = int(parts[0]); vi = vi - 1 if vi > 0 else len(vs) + vi
                    ti = int(parts[1]) if len(parts) > 1 and parts[1] else 0
                    ti = ti - 1 if ti > 0 else (len(vts) + ti if ti < 0 else -1)
                    ni = int(parts[2]) if len(parts) > 2 and parts[2] else 0
                    ni = ni - 1 if ni > 0 else (len(vns) + ni if ni < 0 else -1)
                    key = (vi, ti, ni)
                    if key not in keys:
                        keys[key] = len(out_v); out_v.append(vs[vi])
                        out_t.append(vts[ti] if ti >= 0 else None); out_n.append(vns[ni] if ni >= 0 else None)
                    ids.append(keys[key])
                for k in range(1, len(ids) - 1):
                    tris.append([ids[0], ids[k], ids[k + 1]])
    pos = to_world.transform_affine(np.array(out_v, dtype=np.float64))
    idx = np.array(tris, dtype=np.uint32)
    has_t = all(t is not None for t in out_t) and len(out_t) > 0
    has_n = all(n is not None for n in out_n) and len(out_n) > 0
    uv = None
    if has_t:
        uv = np.array(out_t, dtype=np.float64); uv[:, 1] = 1.0 - uv[:, 1]     # obj.cpp flips v
    nrm = None
    if not face_normals:
        nrm = _xf_normals(to_world, np.array(out_n, dtype=np.float64)) if has_n else compute_vertex_normals(pos.astype(np.float32), idx)
    return MeshData(pos, idx, nrm, uv)
