// lr_scene.cuh -- device-resident scene, LBVH traversal, surface-interaction reconstruction.
//
// HBM layout (all SoA / 128-bit aligned records):
//   nodes  : DNode[n_tris-1]  64 B each = 4 x float4; a node carries BOTH children's boxes so one
//            visit costs four LDG.128 and no dependent child fetch
//   tris   : DTri[n_tris]     48 B each = 3 x float4 (p0,p1,p2 ; w lanes carry mesh / face ids),
//            stored in Morton order (= leaf order) so neighbouring leaves share cache lines
//   grids  : fp32 [z][y][x][c] as the reference (src/volumes/grid.cpp:152-156) + majorant
//            super-grid [mz][my][mx] of local maxima (DDA mode) + fat-voxel copy (DGrid::fat)
#pragma once
#include "lr_math.cuh"
#include "../../include/lr_api.h"

namespace lr {

struct DTri { float4 v0, v1, v2; };   // .w lanes: v0 = global primitive id, v1 = mesh id, v2 = bit 0 null-BSDF mesh, bits 1-2 surface bucket
struct DNode { float4 a, b, c, d; };
// 4-wide quantized node walked by the wavefront TRACE kernels (lr_trace.cuh), 64 B = four 16-byte chunks:
//   chunk 0: origin.xyz (fp32), exponents ex | ey << 8 | ez << 16 (biased, as in an fp32): cell size on axis a = 2^(e_a - 127)
//   chunk 1: child links; >= 0 inner node, < 0 => ~k = k-th triangle of `tris`, 0x7fffffff unused
//   chunk 2: lo.x, lo.y, lo.z, hi.x -- one byte per child, 7-bit grid coordinate with bit 7 SET (0x80 | q), so that a single
//   chunk 3: hi.y, hi.z, -, -        PRMT turns a byte into the float 1 + q/128 (bits 0x3F000000 | byte << 16)
// Plane position = origin + q * cell; lo planes are rounded down and hi planes up on the host, so the quantized box contains the
// exact one.  An unused child has lo = 127, hi = 0 on every axis (never hit).  Nodes are stored in breadth-first order so the top of
// the tree -- what the trace kernels keep in shared memory -- is a prefix of the array.
struct __align__(16) DNodeQ { uint4 c0, c1, c2, c3; };

struct DMesh {
    const float *pos, *nrm, *uv; const uint32_t *idx; const float *area_cdf;
    int bsdf, interior, exterior, emitter;
    uint32_t tri_offset, n_faces; float area; uint32_t flags;      // bit0: medium transition
};

struct DGrid {
    const float *data; float *grad; const float *maj;              // maj: super-grid of local majorants (raw grid units)
    // "fat voxel" copy for 1-channel trilinear clamp grids: record (ix+1, iy+1, iz+1) of (rx+1)(ry+1)(rz+1) holds the eight corner
    // values of the lookup whose base corner is (ix, iy, iz), ix in [-1, rx-1] (indices clamped as the lookup clamps them), in the
    // reference's corner order 000,100,010,110 | 001,101,011,111: one trilinear tap = two LDG.128 from ONE 32-byte sector
    // instead of eight scattered 4-byte loads from four sectors.  Rebuilt from `data` whenever the grid changes.
    const float4 *fat;
    int rx, ry, rz, ch; int mx, my, mz, cell;
    M4 to_local; V3 bmin, bmax; uint32_t filter, wrap; float maxv;
    uint32_t differentiable;
};

// `sealed`: every point of the medium lies inside a closed surface with a non-null BSDF that no emitter shares (lr_backend_api.inl:
// seal_media), so a shadow ray that starts in it always ends on an opaque surface: transmittance exactly 0 (volpath.cpp:528-543)
struct DMedium { lr_medium m; float max_density; uint32_t sealed; };

// `bitmap` texture (src/textures/bitmap.cpp); row-major [h][w][ch] floats in HBM
struct DTex { const float *data; int w, h, ch; uint32_t filter, wrap; float to_uv[6]; };

struct DEmitter {
    lr_emitter e;                     // host pointers inside are NOT valid on the device
    const float *pixels;              // envmap (W+1) x H x 3 with the duplicated column
    const float *warp_data;           // hierarchical warp levels, concatenated
    int n_levels; int level_off[24]; int level_w[24];
    M4 to_world, to_local;
};

#define LR_MAX_BSDFS 32
#define LR_MAX_MEDIA 16
#define LR_MAX_GRIDS 16
#define LR_MAX_EMITTERS 16

// Small, read-only, identical for all threads: passed as a __grid_constant__ kernel parameter
// so every field is a uniform constant-bank load.
struct DScene {
    const DNode *nodes; const DTri *tris; const DMesh *meshes;
    const DNodeQ *nodes4; uint32_t n_nodes4, bvh4_stack;      // 4-wide tree over the same `tris`; worst-case traversal stack depth (entries)
    const DTri *tris_gid;             // triangles by global primitive id (tris is in BVH order)
    const DTex *textures;
    const lr_bsdf *bsdfs; const DMedium *media; const DGrid *grids; const DEmitter *emitters;
    uint32_t n_tris, n_meshes, n_emitters; int env;
    V3 bs_center; float bs_radius;
    lr_sensor sensor; lr_integrator integ;
    M4 cam_to_world, sample_to_camera;
    float gauss_coeff[10];
    uint32_t max_depth; int rr_depth; uint32_t hide_emitters;
    uint32_t prb_handle_null, prb_nee_homogeneous;
    uint32_t any_null_bsdf;           // scene contains index-matched (null) boundaries
};

struct Ray { V3 o, d; float maxt; };
LR_DEV V3 ray_at(const Ray &r, float t) { return fma3(r.d, t, r.o); }

struct Hit { float t; uint32_t prim; float u, v; };
LR_DEV bool hit_valid(const Hit &h) { return h.t != LR_INF; }

// Mesh::moeller_trumbore (include/mitsuba/render/mesh.h:504-526); explicit rounding so the
// result is independent of the compiler's contraction choices (bit-identical to the oracle).
LR_DEV bool moeller_trumbore(const Ray &ray, float maxt, V3 p0, V3 p1, V3 p2, float &t, float &u, float &v) {
    V3 e1 = mk3(__fsub_rn(p1.x, p0.x), __fsub_rn(p1.y, p0.y), __fsub_rn(p1.z, p0.z));
    V3 e2 = mk3(__fsub_rn(p2.x, p0.x), __fsub_rn(p2.y, p0.y), __fsub_rn(p2.z, p0.z));
    V3 pvec = cross(ray.d, e2);
    float inv_det = __frcp_rn(dot(e1, pvec));
    V3 tvec = mk3(__fsub_rn(ray.o.x, p0.x), __fsub_rn(ray.o.y, p0.y), __fsub_rn(ray.o.z, p0.z));
    u = __fmul_rn(dot(tvec, pvec), inv_det);
    bool active = u >= 0.f && u <= 1.f;
    V3 qvec = cross(tvec, e1);
    v = __fmul_rn(dot(ray.d, qvec), inv_det);
    active &= v >= 0.f && __fadd_rn(u, v) <= 1.f;
    t = __fmul_rn(dot(e2, qvec), inv_det);
    active &= t >= 0.f && t <= maxt;
    return active;
}

LR_DEV V3 f4xyz(float4 a) { return mk3(a.x, a.y, a.z); }

// Short-stack LBVH traversal.  `any_hit`: return at the first accepted triangle.
// `skip_null`: ignore triangles of meshes with a null BSDF (opaque-occluder pre-test, DDA mode).
template <bool ANY_HIT, bool SKIP_NULL>
LR_DEV Hit bvh_trace(const DScene &sc, const Ray &ray) {
    Hit h; h.t = LR_INF; h.prim = 0xffffffffu; h.u = h.v = 0.f;
    if (sc.n_tris == 0) return h;
    float tmax = ray.maxt;
    V3 inv = mk3(1.f / ray.d.x, 1.f / ray.d.y, 1.f / ray.d.z);
    auto leaf = [&](uint32_t prim) -> bool {
        const DTri tr = sc.tris[prim];
        if (SKIP_NULL && (__float_as_uint(tr.v2.w) & 1u)) return false;
        float t, u, v;
        if (moeller_trumbore(ray, tmax, f4xyz(tr.v0), f4xyz(tr.v1), f4xyz(tr.v2), t, u, v)) {
            uint32_t gid = __float_as_uint(tr.v0.w);       // global (mesh-order) primitive id: tie-break key
            if (t < h.t || (t == h.t && gid < h.prim)) { h.t = t; h.prim = gid; h.u = u; h.v = v; tmax = t; }
            return true;
        }
        return false;
    };
    if (sc.n_tris == 1) { leaf(0); return h; }
    int stack[64]; int sp = 0; int node = 0;
    while (true) {
        const DNode n = sc.nodes[node];
        // slab test of both children; bounds are padded at build time, far side scaled by 1+4ulp
        float lo0x = (n.a.x - ray.o.x) * inv.x, hi0x = (n.a.w - ray.o.x) * inv.x;
        float lo0y = (n.a.y - ray.o.y) * inv.y, hi0y = (n.b.x - ray.o.y) * inv.y;
        float lo0z = (n.a.z - ray.o.z) * inv.z, hi0z = (n.b.y - ray.o.z) * inv.z;
        float lo1x = (n.b.z - ray.o.x) * inv.x, hi1x = (n.c.y - ray.o.x) * inv.x;
        float lo1y = (n.b.w - ray.o.y) * inv.y, hi1y = (n.c.z - ray.o.y) * inv.y;
        float lo1z = (n.c.x - ray.o.z) * inv.z, hi1z = (n.c.w - ray.o.z) * inv.z;
        // fminf/fmaxf drop NaNs (0 * inf on a slab boundary) like the usual robust formulation
        float n0 = fmaxf(fmaxf(fminf(lo0x, hi0x), fminf(lo0y, hi0y)), fmaxf(fminf(lo0z, hi0z), 0.f));
        float f0 = fminf(fminf(fmaxf(lo0x, hi0x), fmaxf(lo0y, hi0y)), fminf(fmaxf(lo0z, hi0z), tmax)) * 1.0000005f;
        float n1 = fmaxf(fmaxf(fminf(lo1x, hi1x), fminf(lo1y, hi1y)), fmaxf(fminf(lo1z, hi1z), 0.f));
        float f1 = fminf(fminf(fmaxf(lo1x, hi1x), fmaxf(lo1y, hi1y)), fminf(fmaxf(lo1z, hi1z), tmax)) * 1.0000005f;
        bool h0 = n0 <= f0, h1 = n1 <= f1;
        int c0 = __float_as_int(n.d.x), c1 = __float_as_int(n.d.y);
        if (h0 && c0 < 0) { if (leaf((uint32_t) ~c0) && ANY_HIT) return h; h0 = false; }
        if (h1 && c1 < 0) { if (leaf((uint32_t) ~c1) && ANY_HIT) return h; h1 = false; }
        if (h0 && h1) {
            bool swap = n1 < n0;
            int nearc = swap ? c1 : c0, farc = swap ? c0 : c1;
            if (sp < 64) stack[sp++] = farc;
            node = nearc;
        } else if (h0) node = c0;
        else if (h1) node = c1;
        else {
            if (sp == 0) break;
            node = stack[--sp];
        }
    }
    return h;
}

// ---- interactions (include/mitsuba/render/interaction.h) ------------------------------
struct Interaction {
    float t; V3 p, n;
};
LR_DEV V3 offset_p(V3 p, V3 n, V3 d) {                             // interaction.h:165-169
    float mag = (1.f + fmaxf(fabsf(p.x), fmaxf(fabsf(p.y), fabsf(p.z)))) * kRayEps;
    mag = mulsign(mag, dot(n, d));
    return fma3(mk3(mag), n, p);
}
LR_DEV Ray spawn_ray(V3 p, V3 n, V3 d) { Ray r; r.o = offset_p(p, n, d); r.d = d; r.maxt = LR_LARGEST; return r; }
LR_DEV Ray spawn_ray_to(V3 p, V3 n, V3 target) {                   // interaction.h:145-152
    Ray r; r.o = offset_p(p, n, target - p); V3 d = target - r.o; float dist = norm(d);
    r.d = d / dist; r.maxt = dist * (1.f - kShadowEps); return r;
}

struct SurfaceInteraction {
    float t; V3 p, n; Frame sh; V3 wi; V2 uv; int mesh; uint32_t prim;
};
LR_DEV bool si_valid(const SurfaceInteraction &si) { return si.t != LR_INF; }

// Mesh::compute_surface_interaction (src/render/mesh.cpp:1488-1650),
// finalize_surface_interaction (interaction.h:516-536), initialize_sh_frame (:290-300)
LR_DEV SurfaceInteraction compute_si(const DScene &sc, const Ray &ray, const Hit &h, const DTri *tri_by_gid) {
    SurfaceInteraction si;
    if (!hit_valid(h)) {
        si.t = LR_INF; si.wi = -ray.d; si.mesh = -1; si.prim = 0; si.p = mk3(0.f); si.n = mk3(0.f);
        si.sh.n = si.sh.s = si.sh.t = mk3(0.f); si.uv = mk2(0.f, 0.f); return si;
    }
    const DTri tr = tri_by_gid[h.prim];
    uint32_t mesh_id = __float_as_uint(tr.v1.w);
    const DMesh &me = sc.meshes[mesh_id];
    uint32_t face = h.prim - me.tri_offset;
    uint32_t i0 = me.idx[3 * face], i1 = me.idx[3 * face + 1], i2 = me.idx[3 * face + 2];
    V3 p0 = f4xyz(tr.v0), p1 = f4xyz(tr.v1), p2 = f4xyz(tr.v2);
    float b1 = h.u, b2 = h.v, b0 = __fsub_rn(__fsub_rn(1.f, b1), b2);
    si.p = fma3(p0, b0, fma3(p1, b1, mk3(__fmul_rn(p2.x, b2), __fmul_rn(p2.y, b2), __fmul_rn(p2.z, b2))));
    si.t = h.t; si.mesh = (int) mesh_id; si.prim = h.prim;
    V3 dp0 = p1 - p0, dp1 = p2 - p0;
    si.n = normalize(cross(dp0, dp1));
    si.uv = mk2(b1, b2);
    V3 dp_du, dp_dv; coordinate_system(si.n, dp_du, dp_dv);
    if (me.uv) {
        V2 uv0 = mk2(me.uv[2 * i0], me.uv[2 * i0 + 1]), uv1 = mk2(me.uv[2 * i1], me.uv[2 * i1 + 1]), uv2 = mk2(me.uv[2 * i2], me.uv[2 * i2 + 1]);
        si.uv = mk2(fmaf(uv2.x, b2, fmaf(uv1.x, b1, __fmul_rn(uv0.x, b0))), fmaf(uv2.y, b2, fmaf(uv1.y, b1, __fmul_rn(uv0.y, b0))));
        V2 duv0 = mk2(uv1.x - uv0.x, uv1.y - uv0.y), duv1 = mk2(uv2.x - uv0.x, uv2.y - uv0.y);
        float det = fmaf(duv0.x, duv1.y, -__fmul_rn(duv0.y, duv1.x));
        if (det != 0.f) {
            float inv_det = 1.f / det;
            dp_du = mk3(fmaf(duv1.y, dp0.x, -__fmul_rn(duv0.y, dp1.x)), fmaf(duv1.y, dp0.y, -__fmul_rn(duv0.y, dp1.y)),
                        fmaf(duv1.y, dp0.z, -__fmul_rn(duv0.y, dp1.z))) * inv_det;
        }
    }
    if (me.nrm) {
        V3 n0 = mk3(me.nrm[3 * i0], me.nrm[3 * i0 + 1], me.nrm[3 * i0 + 2]), n1 = mk3(me.nrm[3 * i1], me.nrm[3 * i1 + 1], me.nrm[3 * i1 + 2]),
           n2 = mk3(me.nrm[3 * i2], me.nrm[3 * i2 + 1], me.nrm[3 * i2 + 2]);
        V3 n = fma3(n2, b2, fma3(n1, b1, mk3(__fmul_rn(n0.x, b0), __fmul_rn(n0.y, b0), __fmul_rn(n0.z, b0))));
        si.sh.n = n * (1.f / sqrtf(sqnorm(n)));
    } else si.sh.n = si.n;
    if (me.flags & 2u) { si.n = -si.n; si.sh.n = -si.sh.n; }         // flip_normals (mesh.cpp:1652-1655)
    si.sh.s = normalize(fma3(si.sh.n, -dot(si.sh.n, dp_du), dp_du));
    if (dp_du.x == 0.f && dp_du.y == 0.f && dp_du.z == 0.f) { V3 t_; coordinate_system(si.sh.n, si.sh.s, t_); }
    si.sh.t = cross(si.sh.n, si.sh.s);
    si.wi = si.sh.to_local(-ray.d);
    return si;
}

} // namespace lr
