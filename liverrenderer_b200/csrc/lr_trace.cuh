// lr_trace.cuh -- ray queries of the wavefront integrators as their own persistent stage (Scene::ray_intersect / ray_test,
// src/render/scene.cpp:204-248; Mesh::moeller_trumbore, include/mitsuba/render/mesh.h:504-526).
//
// A trace kernel owns nothing but rays: 32 B in (origin + maxt, direction), 16 B out (t, prim, u, v).  Each thread walks the
// 4-wide tree (DNode4, lr_scene.cuh) with its traversal stack in SHARED memory ([depth][thread]: conflict-free, no local
// memory, no spills); the stack depth is the exact worst case of the tree, computed by the host builder (bvh4_stack), and
// the kernel is launched with that much dynamic shared memory.  The kernel is persistent: a lane whose ray has finished
// parks, and as soon as LR_TRACE_REFILL lanes of the warp are parked they write their hits out, route their paths to the
// next stage's queue and pull fresh rays (one atomicAdd on the queue cursor per refill), so the node loop stays populated
// although rays of one warp need very different numbers of visits.
//
// Hits do not depend on the tree: every candidate triangle goes through the same exactly-rounded Moeller-Trumbore test,
// ties go to the lower global primitive id, and leaf boxes are padded (lr_bvh_host.inl), so this traversal, the two-wide
// bvh_trace() of the per-lane kernels and the oracle's brute force return identical (t, prim, u, v).
#pragma once
#include "lr_scene.cuh"

#ifndef LR_TRACE_REFILL
#define LR_TRACE_REFILL 12            // refill once at least this many lanes of the warp are parked
#endif

namespace lr {

// Per-thread traversal state of one ray.  `node`: next node to visit, -1 = finished (parked), -2 = no ray.
struct TraceRay {
    V3 o, d, inv, ood; float tmax; Hit h; int node; int sp;
};

LR_DEV void trace_begin(const DScene &sc, TraceRay &r, V3 o, V3 d, float maxt) {
    r.o = o; r.d = d; r.tmax = maxt;
    r.inv = mk3(1.f / d.x, 1.f / d.y, 1.f / d.z);
    r.ood = mk3(-o.x * r.inv.x, -o.y * r.inv.y, -o.z * r.inv.z);
    r.h.t = LR_INF; r.h.prim = 0xffffffffu; r.h.u = r.h.v = 0.f;
    r.sp = 0; r.node = sc.n_tris ? 0 : -1;
}

// slab test of child `i` of a node: entry distance, or +inf when the ray misses the (padded) box within [0, tmax].
// fminf / fmaxf drop the NaNs of 0 * inf on a slab boundary and of an axis-parallel ray (then that axis does not cull:
// conservative).  The far side is widened by 4 ulp like the two-wide traversal does.
#define LR_SLAB(i, C)                                                                                                   \
    {                                                                                                                   \
        float ax = fmaf(n_lox.C, r.inv.x, r.ood.x), bx = fmaf(n_hix.C, r.inv.x, r.ood.x);                               \
        float ay = fmaf(n_loy.C, r.inv.y, r.ood.y), by = fmaf(n_hiy.C, r.inv.y, r.ood.y);                               \
        float az = fmaf(n_loz.C, r.inv.z, r.ood.z), bz = fmaf(n_hiz.C, r.inv.z, r.ood.z);                               \
        float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.f));                               \
        float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), r.tmax)) * 1.0000005f;               \
        t[i] = tn <= tf ? tn : LR_INF;                                                                                  \
    }

// One node visit.  ANY_HIT: stop at the first accepted triangle (occlusion query); SKIP_NULL: ignore triangles of
// index-matched (null-BSDF) meshes.  `stack` points at this thread's column of the shared stack, stride `stride` words.
template <bool ANY_HIT, bool SKIP_NULL>
LR_DEV void trace_step(const DScene &sc, TraceRay &r, int *stack, const int stride) {
    const DNode4 *nd = sc.nodes4 + r.node;
    const float4 n_lox = __ldg(&nd->lox), n_loy = __ldg(&nd->loy), n_loz = __ldg(&nd->loz);
    const float4 n_hix = __ldg(&nd->hix), n_hiy = __ldg(&nd->hiy), n_hiz = __ldg(&nd->hiz);
    const int4 ch = __ldg(&nd->child);
    float t[4]; int c[4] = { ch.x, ch.y, ch.z, ch.w };
    LR_SLAB(0, x) LR_SLAB(1, y) LR_SLAB(2, z) LR_SLAB(3, w)
    // leaves first: a hit shortens the ray before the inner children are ordered
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (t[i] != LR_INF && c[i] < 0) {
            const DTri tr = sc.tris[~c[i]];
            if (!(SKIP_NULL && (__float_as_uint(tr.v2.w) & 1u))) {
                Ray ray; ray.o = r.o; ray.d = r.d; ray.maxt = r.tmax;
                float tt, u, v;
                if (moeller_trumbore(ray, r.tmax, f4xyz(tr.v0), f4xyz(tr.v1), f4xyz(tr.v2), tt, u, v)) {
                    const uint32_t gid = __float_as_uint(tr.v0.w);
                    if (tt < r.h.t || (tt == r.h.t && gid < r.h.prim)) { r.h.t = tt; r.h.prim = gid; r.h.u = u; r.h.v = v; r.tmax = tt; }
                    if (ANY_HIT) { r.node = -1; return; }
                }
            }
            t[i] = LR_INF;
        }
    }
    // inner children whose entry lies beyond a hit just found are dropped
#pragma unroll
    for (int i = 0; i < 4; ++i) if (t[i] > r.tmax) t[i] = LR_INF;
    if (!ANY_HIT) {
        // sort the (entry distance, child) pairs: 5-comparator network, nearest first
#define LR_CSWAP(a, b) { const bool s = t[b] < t[a]; const float tlo = s ? t[b] : t[a], thi = s ? t[a] : t[b]; \
                         const int clo = s ? c[b] : c[a], chi = s ? c[a] : c[b]; t[a] = tlo; t[b] = thi; c[a] = clo; c[b] = chi; }
        LR_CSWAP(0, 1) LR_CSWAP(2, 3) LR_CSWAP(0, 2) LR_CSWAP(1, 3) LR_CSWAP(1, 2)
#undef LR_CSWAP
        if (t[3] != LR_INF) { stack[r.sp * stride] = c[3]; ++r.sp; }
        if (t[2] != LR_INF) { stack[r.sp * stride] = c[2]; ++r.sp; }
        if (t[1] != LR_INF) { stack[r.sp * stride] = c[1]; ++r.sp; }
        if (t[0] != LR_INF) { r.node = c[0]; return; }
    } else {
        int next = -1;
#pragma unroll
        for (int i = 0; i < 4; ++i) if (t[i] != LR_INF) { if (next >= 0) { stack[r.sp * stride] = next; ++r.sp; } next = c[i]; }
        if (next >= 0) { r.node = next; return; }
    }
    if (r.sp == 0) { r.node = -1; return; }
    --r.sp; r.node = stack[r.sp * stride];
}
#undef LR_SLAB

} // namespace lr
