// lr_trace.cuh -- ray queries of the wavefront integrators as their own persistent stage (Scene::ray_intersect / ray_test,
// src/render/scene.cpp:204-248; Mesh::moeller_trumbore, include/mitsuba/render/mesh.h:504-526).
//
// A trace kernel owns nothing but rays: 32 B in (origin + maxt, direction), 16 B out (t, prim, u, v).  Each thread walks a
// 4-wide tree of 64-byte quantized nodes (DNodeQ, lr_scene.cuh).  What bounds incoherent traversal on this machine is the L1
// data pipe -- one wavefront per distinct line per load instruction (ncu: l1tex__data_pipe_lsu_wavefronts at 93 % of the
// cycles with 128-byte fp32 nodes in global memory, profiles/r2_trace_kernel.txt) -- so the top of the tree (all of it for the
// liver scenes) is copied into SHARED memory by every block, a visit is four LDS.128 on bank-swizzled chunks, and the
// traversal stack lives in shared memory too ([depth][thread], 16-bit entries when the tree is small; entries beyond
// LR_TRACE_SD go to a global scratch column).  The kernel is persistent: a lane whose ray has finished parks, and as soon as
// LR_TRACE_REFILL lanes of the warp are parked they write their hits out, route their paths to the next stage's queue and pull
// fresh rays, so the node loop stays populated although rays of one warp need very different numbers of visits.
//
// Hits do not depend on the tree: every candidate triangle goes through the same exactly-rounded Moeller-Trumbore test,
// ties go to the lower global primitive id, and boxes are padded / rounded outward (lr_bvh_host.inl), so this traversal, the
// two-wide bvh_trace() of the per-lane kernels and the oracle's brute force return identical (t, prim, u, v).
#pragma once
#include "lr_scene.cuh"

#ifndef LR_TRACE_REFILL
#define LR_TRACE_REFILL 12            // refill once at least this many lanes of the warp are parked
#endif
#ifndef LR_TRACE_LEAF_BATCH
#define LR_TRACE_LEAF_BATCH 10        // triangle tests run once this many lanes of the warp hold a triangle (or no lane has a node to visit)
#endif
#ifndef LR_TRACE_SD
#define LR_TRACE_SD 16                // stack entries per lane kept in shared memory
#endif
#ifndef LR_TRACE_THREADS
#define LR_TRACE_THREADS 1024         // threads per block, one block per SM
#endif

namespace lr {

// Per-thread traversal state of one ray.  `node`: what the lane does next -- an inner node to visit (>= 0, below TR_DONE),
// a triangle to test (< 0: ~k = k-th triangle of `tris`), TR_DONE = finished (parked), TR_IDLE = no ray.
#define TR_DONE 0x7ffffffe
#define TR_IDLE 0x7fffffff
struct TraceRay {
    V3 o, d, inv, ood; float tmax; Hit h; int node; int sp;
};
LR_DEV bool trace_walking(const TraceRay &r) { return r.node < TR_DONE; }

LR_DEV void trace_begin(const DScene &sc, TraceRay &r, V3 o, V3 d, float maxt) {
    r.o = o; r.d = d; r.tmax = maxt;
    // 1 / d, kept FINITE: the quantized slab test forms t = f * A + B with A ~ inv and B ~ -o * inv - A, and with inv = inf (a direction
    // component that is exactly 0: an isotropic phase sample at u = 0 gives d = (0, 0, 1), one ray in ~1e7) every plane of that axis
    // evaluates to inf - inf = NaN, fmaxf / fminf drop it, and the axis stops culling: such a ray visited every box that overlaps it along
    // its one live axis -- thousands of nodes, one lane wide, 15 ms at the tail of a launch on the 20 k-triangle sphere.  With |inv|
    // capped the planes are huge finite numbers of the right sign and the slabs cull exactly as they do for fp32 boxes.
    const float kInvMax = 1e25f;
    r.inv = mk3(copysignf(fminf(fabsf(1.f / d.x), kInvMax), d.x), copysignf(fminf(fabsf(1.f / d.y), kInvMax), d.y), copysignf(fminf(fabsf(1.f / d.z), kInvMax), d.z));
    r.ood = mk3(-o.x * r.inv.x, -o.y * r.inv.y, -o.z * r.inv.z);
    r.h.t = LR_INF; r.h.prim = 0xffffffffu; r.h.u = r.h.v = 0.f;
    r.sp = 0; r.node = sc.n_tris ? 0 : TR_DONE;
    // A ray with a non-finite origin or direction (a degenerate BSDF / phase sample, about one path in 1e7) hits nothing in any
    // traversal here -- every comparison of its slab test is false in the two-wide walk.  In the wide walk fmaxf / fminf DROP the NaNs,
    // no box would cull, and the ray would visit the whole tree one lane wide (20 k triangles: tens of milliseconds at the tail of a
    // launch): it is finished as a miss at once.
    // The same goes for a null direction (inv = inf on every axis => every slab interval is inf - inf).
    if (!isfinite((o.x + o.y + o.z) + (d.x + d.y + d.z)) || maxt != maxt || (d.x == 0.f && d.y == 0.f && d.z == 0.f)) r.node = TR_DONE;
}

// The traversal stack: the first `sd` entries of a lane live in shared memory (column `stack`, stride `stride` elements),
// deeper ones -- rare: the worst case of the tree is far above what rays reach -- in a global scratch column `ovf`.
template <typename STK>
struct TraceStack { STK *stack; int *ovf; int stride, ovf_stride, sd; };
template <typename STK>
LR_DEV void trace_push(TraceRay &r, const TraceStack<STK> &st, int v) {
    if (r.sp < st.sd) st.stack[r.sp * st.stride] = (STK) v; else st.ovf[(size_t) (r.sp - st.sd) * st.ovf_stride] = v;
    ++r.sp;
}
template <typename STK>
LR_DEV void trace_pop(TraceRay &r, const TraceStack<STK> &st) {
    if (r.sp == 0) { r.node = TR_DONE; return; }
    --r.sp; r.node = r.sp < st.sd ? (int) st.stack[r.sp * st.stride] : st.ovf[(size_t) (r.sp - st.sd) * st.ovf_stride];
}

// Visit of an inner node (DNodeQ, lr_scene.cuh): the hit children -- inner nodes and triangles alike -- go onto the stack
// farthest first and the nearest becomes the lane's next item (occlusion queries use the same order: near boxes first).
// Triangles are NOT tested here: lanes that reach a triangle wait until enough lanes of the warp hold one (trace kernel loop), so the Moeller-Trumbore code runs with many
// lanes instead of 2-3.  Nodes below `n_smem` are read from the shared-memory copy `s_nodes` (16-byte chunks rotated by
// node >> 1 so that lanes at different nodes spread over all bank groups), the others from global memory.
// Slab test per axis: plane = origin + q * cell, t = (plane - o) / d = f * A + B with f = 1 + q/128 (one PRMT per byte),
// A = 128 * cell / d, B = (origin - o) / d - A.  Near / far planes are picked by the sign of d, so an unused child
// (lo = 127 > hi = 0) fails, and fmaxf / fminf drop the NaNs of an axis-parallel ray (that axis then does not cull:
// conservative).  The far side is widened by 4 ulp like the two-wide traversal does.
template <typename STK>
LR_DEV void trace_visit(const DScene &sc, TraceRay &r, const uint4 *s_nodes, const uint32_t n_smem, const TraceStack<STK> &st) {
    const uint32_t n = (uint32_t) r.node; uint4 c0, c1, c2, c3;
    if (n < n_smem) {
        const uint4 *b = s_nodes + n * 4u; const uint32_t sw = n >> 1;
        c0 = b[sw & 3u]; c1 = b[(sw + 1u) & 3u]; c2 = b[(sw + 2u) & 3u]; c3 = b[(sw + 3u) & 3u];
    } else {
        const DNodeQ *g = sc.nodes4 + n; c0 = __ldg(&g->c0); c1 = __ldg(&g->c1); c2 = __ldg(&g->c2); c3 = __ldg(&g->c3);
    }
    // 128 * cell * inv: the biased exponent byte + 7
    const float Ax = __uint_as_float(((c0.w & 0xffu) + 7u) << 23) * r.inv.x, Ay = __uint_as_float((((c0.w >> 8) & 0xffu) + 7u) << 23) * r.inv.y,
                Az = __uint_as_float((((c0.w >> 16) & 0xffu) + 7u) << 23) * r.inv.z;
    const float Bx = fmaf(__uint_as_float(c0.x), r.inv.x, r.ood.x) - Ax, By = fmaf(__uint_as_float(c0.y), r.inv.y, r.ood.y) - Ay,
                Bz = fmaf(__uint_as_float(c0.z), r.inv.z, r.ood.z) - Az;
    const bool nx = r.inv.x < 0.f, ny = r.inv.y < 0.f, nz = r.inv.z < 0.f;
    const uint32_t wxn = nx ? c2.w : c2.x, wxf = nx ? c2.x : c2.w, wyn = ny ? c3.x : c2.y, wyf = ny ? c2.y : c3.x, wzn = nz ? c3.y : c2.z, wzf = nz ? c2.z : c3.y;
    float t[4]; int c[4] = { (int) c1.x, (int) c1.y, (int) c1.z, (int) c1.w };
#define LR_PL(w, i) __uint_as_float(__byte_perm(w, 0x3F000000u, 0x7044u | ((i) << 8)))
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float tn = fmaxf(fmaxf(fmaf(LR_PL(wxn, i), Ax, Bx), fmaf(LR_PL(wyn, i), Ay, By)), fmaxf(fmaf(LR_PL(wzn, i), Az, Bz), 0.f));
        const float tf = fminf(fminf(fmaf(LR_PL(wxf, i), Ax, Bx), fmaf(LR_PL(wyf, i), Ay, By)), fminf(fmaf(LR_PL(wzf, i), Az, Bz), r.tmax)) * 1.0000005f;
        t[i] = (tn <= tf && c[i] != 0x7fffffff) ? tn : LR_INF;
    }
#undef LR_PL
    // sort the (entry distance, child) pairs: 5-comparator network, nearest first (misses last)
#define LR_CSWAP(a, b) { const bool s = t[b] < t[a]; const float tlo = s ? t[b] : t[a], thi = s ? t[a] : t[b]; \
                         const int clo = s ? c[b] : c[a], chi = s ? c[a] : c[b]; t[a] = tlo; t[b] = thi; c[a] = clo; c[b] = chi; }
    LR_CSWAP(0, 1) LR_CSWAP(2, 3) LR_CSWAP(0, 2) LR_CSWAP(1, 3) LR_CSWAP(1, 2)
#undef LR_CSWAP
    if (t[3] != LR_INF) trace_push(r, st, c[3]);
    if (t[2] != LR_INF) trace_push(r, st, c[2]);
    if (t[1] != LR_INF) trace_push(r, st, c[1]);
    if (t[0] != LR_INF) { r.node = c[0]; return; }
    trace_pop(r, st);
}

// Test of the triangle the lane holds.  ANY_HIT: the first accepted triangle ends the query (occlusion); SKIP_NULL: triangles of
// index-matched (null-BSDF) meshes are ignored.
template <bool ANY_HIT, bool SKIP_NULL, typename STK>
LR_DEV void trace_leaf(const DScene &sc, TraceRay &r, const TraceStack<STK> &st) {
    const DTri tr = sc.tris[~r.node];
    if (!(SKIP_NULL && (__float_as_uint(tr.v2.w) & 1u))) {
        Ray ray; ray.o = r.o; ray.d = r.d; ray.maxt = r.tmax;
        float tt, u, v;
        if (moeller_trumbore(ray, r.tmax, f4xyz(tr.v0), f4xyz(tr.v1), f4xyz(tr.v2), tt, u, v)) {
            const uint32_t gid = __float_as_uint(tr.v0.w);
            if (tt < r.h.t || (tt == r.h.t && gid < r.h.prim)) { r.h.t = tt; r.h.prim = gid; r.h.u = u; r.h.v = v; r.tmax = tt; }
            if (ANY_HIT) { r.node = TR_DONE; return; }
        }
    }
    trace_pop(r, st);
}

} // namespace lr
