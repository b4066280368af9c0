// lr_wavefront.cuh -- wavefront `volpath` (reference: src/integrators/volpath.cpp:93-554), the production forward path.
// The per-lane loop of sample_volpath() (lr_integrators.cuh) is cut where lanes diverge; the pieces are stage kernels over
// index queues, and ray queries are a stage of their own (lr_trace.cuh):
//
//   GEN       lane hand-out, TEA/PCG32 seeding, camera ray, primary hit.  A primary ray that leaves the scene is finished
//             right here (environment lookup + film splat, no path state is ever written for it)
//   TRACK     free-flight sampling in the current medium (null collisions stay in registers) until an EVENT:
//             real scatter -> scatter queue | boundary reached -> surface bucket / escape queue | depth limit -> film
//   SHADE_A   scatter: emitter sample + phase value; surface: emitter hit + MIS, emitter sample + BSDF value.  Writes the
//             shadow ray, pushes the slot to the shadow queue
//   ESCAPE    rays that left the scene: environment radiance + MIS, film splat
//   TRACE_S   occlusion of the shadow rays (persistent, 4-wide BVH); unoccluded segments outside any medium are added to
//             the path's radiance at once, the others go to the NEE queue
//   NEE       ratio-tracking transmittance walk along the shadow segment (volpath.cpp:454-551)
//   SHADE_B   phase / BSDF sampling, medium transition, then the russian-roulette + depth test that opens the reference's
//             NEXT loop iteration (volpath.cpp:160-178; no draw lies between the two, so the stream order is the
//             reference's): a path that dies never costs a ray.  Survivors go to the trace queue
//   TRACE_C   closest hit of the new rays (persistent, 4-wide BVH); routes each path by what comes next:
//             inside a medium -> TRACK queue, surface -> bucket of its BSDF class, nothing hit -> escape queue
//
// Path state lives in SoA planes of 16-byte records indexed by slot (WVPlanes): a stage touches only the planes it needs
// (TRACE: 32 B in / 16 B out), and consecutive queue entries -- queues are compacted with warp ballots in lane order --
// share sectors.  Every lane keeps its own PCG32 stream (keyed by the GLOBAL lane index) and draws in exactly the order
// of the reference, so a wavefront render is lane-for-lane identical to the per-lane kernel and to the oracle.
#pragma once
#include "lr_wf_common.cuh"
#include "lr_trace.cuh"
#include "lr_tma.cuh"
#include <cuda_pipeline.h>

namespace lr {

// Path state, indexed by slot.  Fields that are always touched together share one 32-byte record = one DRAM sector (after the
// first bounce the queues address slots in no particular order, so every access costs whole sectors; 16-byte SoA planes would use
// half of each):
//   h0  { PCG32 state, inc | throughput, depth }             every stage that draws or scales
//   h1  { ray.o, ray.maxt  | ray.d, flags }                  the ray: all TRACE needs
//   h2  { radiance so far, eta | last_scatter_p, last_scatter_direction_pdf }   emitter hits, MIS, russian roulette beyond rr_depth
//   hit { t, prim, u, v }   lane { lane lo/hi, film sample position }           16-byte records
//   n0  { shadow ray origin, max_dist | shadow ray direction, ds.dist }         NEE (volpath.cpp:400-554)
//   n1  { throughput * (phase | bsdf) value * MIS weight * emitter value, walk medium | closest hit along the shadow ray }
//   adj { dLoss / dRadiance of the lane }                                      16 bytes, touched by the prbvolpath adjoint only
struct WVPlanes {
    float4 *h0, *h1, *h2, *hit_, *lane_, *n0, *n1, *adj_;
    LR_DEV ulonglong2 &rng(uint32_t s) const { return reinterpret_cast<ulonglong2 *>(h0)[2 * (size_t) s]; }
    LR_DEV float4 &thr(uint32_t s) const { return h0[2 * (size_t) s + 1]; }
    LR_DEV float4 &o(uint32_t s) const { return h1[2 * (size_t) s]; }
    LR_DEV float4 &d(uint32_t s) const { return h1[2 * (size_t) s + 1]; }
    LR_DEV float4 &res(uint32_t s) const { return h2[2 * (size_t) s]; }
    LR_DEV float4 &lsp(uint32_t s) const { return h2[2 * (size_t) s + 1]; }
    LR_DEV float4 &hit(uint32_t s) const { return hit_[s]; }
    LR_DEV float4 &lane(uint32_t s) const { return lane_[s]; }
    LR_DEV float4 &no(uint32_t s) const { return n0[2 * (size_t) s]; }
    LR_DEV float4 &nd(uint32_t s) const { return n0[2 * (size_t) s + 1]; }
    LR_DEV float4 &nw(uint32_t s) const { return n1[2 * (size_t) s]; }
    LR_DEV float4 &nhit(uint32_t s) const { return n1[2 * (size_t) s + 1]; }
    LR_DEV float4 &adj(uint32_t s) const { return adj_[s]; }          // prbvolpath adjoint: delta L of the lane (lr_wavefront_prb.cuh)
};
#define WV_N_PLANES 13                  // 16-byte units per slot

struct WVQueues {                      // device-side counters of one pool
    uint32_t n_trace, n_track, n_scatter, n_escape, n_shadow, n_nee, n_regen;
    uint32_t n_surf[WF_N_SURF]; uint32_t cursor[3];                    // work cursors of TRACE_C, TRACK, TRACE_S
    uint32_t n_alloc, live, cursor_nee, pad[2];                        // cursor_nee: work cursor of the persistent NEE walk (lr_wavefront_mis.cuh)
};
#define WV_N_QUEUES (7 + WF_N_SURF)

struct WVArgs {
    WVPlanes p; WVQueues *q;
    uint32_t *q_trace, *q_track, *q_scatter, *q_escape, *q_shadow, *q_nee, *q_regen, *q_surf[WF_N_SURF];
    uint32_t cap, slack;                                               // slots; GEN stops handing out lanes at cap - slack
    float *film; float *lanes_out;
    uint64_t lane_begin, lane_end; unsigned long long *lane_next;      // lane cursor shared by all batches of one render call
    uint32_t seed, spp_per_pass, n_passes;
    unsigned long long *stats; LaneMap map;
    float4 *mis;                                                       // volpathmis: eight 16-byte records per slot (lr_wavefront_mis.cuh), else nullptr
};

LR_DEV Sampler wv_rng(const WVPlanes &p, uint32_t s) { const ulonglong2 r = p.rng(s); Sampler m; m.state = r.x; m.inc = r.y; return m; }
LR_DEV void wv_rng_store(const WVPlanes &p, uint32_t s, const Sampler &m) { p.rng(s) = make_ulonglong2(m.state, m.inc); }
LR_DEV uint32_t wv_bucket(const DScene &sc, const Hit &h) { return (__float_as_uint(sc.tris_gid[h.prim].v2.w) >> 1) & 3u; }

LR_DEV bool wv_finite(V3 v) { return isfinite(v.x + v.y + v.z); }

// volpath.cpp:160-178: the russian-roulette + depth test at the top of a loop iteration
// (a throughput that is no longer finite -- 0 * inf after an underflowed pdf, about one path in 1e6 on the nested-shell scene -- ends the
// path: the reference would carry the NaN into the pixel)
LR_DEV bool wv_rr(const DScene &sc, Sampler &smp, V3 &thr, float eta, uint32_t depth) {
    bool active = any_nonzero(thr) && wv_finite(thr);
    const float q = fminf(hmax(thr) * sqr(eta), .95f);
    const bool perform_rr = depth > (uint32_t) sc.rr_depth;
    if (active) { const float u = smp.next_1d(); active &= (u < q) || !perform_rr; }
    if (perform_rr) thr *= 1.f / q;
    active &= depth < sc.max_depth;
    return active;
}

// End of a path: ImageBlock::put or the per-lane output.  Call where the warp is converged; `pred`: this lane finishes.
// Box filter: lanes of one pixel are neighbours in the queues (lane = pixel * spp + sample), so a warp that finishes samples of a
// single pixel adds them up with shuffles and issues one set of atomics instead of 32 to the same address.
template <int C, bool TO_FILM>
LR_DEV void wv_finish(const DScene &sc, const WVArgs &a, bool pred, float4 lane_pos, V3 L, uint32_t flags) {
    const bool valid = (flags & WF_VALID_RAY) != 0;
    if (!wv_finite(L)) L = mk3(0.f);              // a non-finite path (see wv_rr) contributes nothing instead of poisoning its pixel
    if (!TO_FILM) {
        if (pred) {
            const uint64_t lane = ((uint64_t) __float_as_uint(lane_pos.y) << 32) | __float_as_uint(lane_pos.x);
            const uint32_t pass = flags >> WF_PASS_SHIFT;
            reinterpret_cast<float4 *>(a.lanes_out)[(lane_local(a.map, lane) - a.lane_begin) * a.n_passes + pass] = make_float4(L.x, L.y, L.z, valid ? 1.f : 0.f);
        }
        return;
    }
    float aovs[C]; aovs[0] = L.x; aovs[1] = L.y; aovs[2] = L.z;
    if (C == 5) { aovs[3] = valid ? 1.f : 0.f; aovs[4] = 1.f; } else aovs[3] = 1.f;
    const unsigned m = __ballot_sync(0xffffffffu, pred);
    if (m == 0) return;
    if (sc.sensor.rfilter == LR_RFILTER_BOX && __popc(m) > 1) {
        const int W = (int) sc.sensor.film_width, H = (int) sc.sensor.film_height;
        const int x = (int) floorf(lane_pos.z), y = (int) floorf(lane_pos.w);
        const int pix = (pred && x >= 0 && y >= 0 && x < W && y < H) ? y * W + x : -1;
        const int leader = __ffs(m) - 1, pix0 = __shfl_sync(0xffffffffu, pix, leader);
        if (__all_sync(0xffffffffu, !pred || pix == pix0)) {
            if (pix0 < 0) return;
#pragma unroll
            for (int k = 0; k < C; ++k) {
                float v = pred ? aovs[k] : 0.f;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if ((int) (threadIdx.x & 31u) == leader) atomicAdd(a.film + (size_t) pix0 * C + k, v);
            }
            return;
        }
    }
    if (pred) film_put<C>(sc, a.film, mk2(lane_pos.z, lane_pos.w), aovs, a.spp_per_pass >= 4);
}

// a finished path whose lane has further passes to render (integrator.cpp:347-356: the stream simply continues)
LR_DEV bool wv_more_passes(const WVArgs &a, uint32_t flags) { return a.n_passes > 1 && (flags >> WF_PASS_SHIFT) + 1 < a.n_passes; }

// where a path with a fresh closest hit goes next.  One round trip for all five queues: lane k of the warp reserves the
// entries of queue k (TRACK, ESCAPE, surface buckets), the bases travel back by shuffle.
LR_DEV void wv_route(const DScene &sc, const WVArgs &a, uint32_t slot, bool pred, uint32_t flags, const Hit &h) {
    uint32_t dest = 0xffu;                                                  // 0 TRACK, 1 ESCAPE, 2.. surface bucket
    if (pred) dest = wf_medium(flags) >= 0 ? 0u : (!hit_valid(h) ? 1u : 2u + wv_bucket(sc, h));
    const unsigned lane = threadIdx.x & 31u;
    unsigned mine = 0, m[2 + WF_N_SURF];
#pragma unroll
    for (uint32_t k = 0; k < 2 + WF_N_SURF; ++k) { m[k] = __ballot_sync(0xffffffffu, dest == k); if (dest == k) mine = m[k]; }
    uint32_t base = 0;
    if (lane < 2 + WF_N_SURF) {
        unsigned mk = 0;
#pragma unroll
        for (uint32_t k = 0; k < 2 + WF_N_SURF; ++k) if (lane == k) mk = m[k];
        uint32_t *cnt = lane == 0 ? &a.q->n_track : (lane == 1 ? &a.q->n_escape : &a.q->n_surf[lane - 2]);
        if (mk) base = atomicAdd(cnt, (uint32_t) __popc(mk));
    }
    base = __shfl_sync(0xffffffffu, base, dest & 31u);
    if (pred) {
        uint32_t *q = dest == 0 ? a.q_track : (dest == 1 ? a.q_escape : a.q_surf[dest - 2]);
        q[base + __popc(mine & ((1u << lane) - 1u))] = slot;
    }
}

// ------------------------------------------------------------------------------------------
// GEN (integrator.cpp:321-356, perspective.cpp:239-279, volpath.cpp:93-178): starts lanes.  `own_slot`: the lane already owns
// `slot` (a later pass of a multi-pass render); otherwise a slot is allocated only if the path survives its first segment.
// ------------------------------------------------------------------------------------------
template <int C, bool TO_FILM>
LR_DEV void wv_start_paths(const DScene &sc, const WVArgs &a, bool have, uint64_t lane, uint32_t pass, Sampler smp, bool own_slot, uint32_t slot,
                           Counters &cnt, unsigned long long &nsamples) {
    LaneSetup L; L.smp = smp; L.pos = mk2(0.f, 0.f); L.sample_pos = L.pos; L.ray.o = L.ray.d = mk3(0.f); L.ray.maxt = 0.f;
    if (have) lane_begin_sample(sc, L, lane, a.spp_per_pass);
    const bool box = sc.sensor.rfilter == LR_RFILTER_BOX;
    bool working = have, keep = false; Hit h; h.t = LR_INF; h.prim = 0xffffffffu; h.u = h.v = 0.f; uint32_t f = 0; V3 thr = mk3(1.f);
    while (__any_sync(0xffffffffu, working)) {
        bool fin = false; V3 Lr = mk3(0.f);
        if (working) {
            lane_next_ray(sc, L);
            const uint32_t channel = (uint32_t) fminf(L.smp.next_1d() * 3.f, 2.f);           // volpath.cpp:117-120
            f = channel | WF_HAS_LANE | (pass << WF_PASS_SHIFT);
            if (!sc.hide_emitters && sc.env >= 0) f |= WF_VALID_RAY;                          // :103
            if (!sc.hide_emitters) f |= WF_SPEC_CHAIN;                                        // :114
            f = wf_set_medium(f, sc.sensor.medium);
            nsamples++; thr = mk3(1.f);
            const bool alive = wv_rr(sc, L.smp, thr, 1.f, 0u);                                // :160-178 of the first iteration
            if (alive) { cnt.rays++; h = bvh_trace<false, false>(sc, L.ray); }
            if (!alive) fin = true;
            else if (!hit_valid(h) && sc.sensor.medium < 0) {                                 // left the scene: :326-345 with an invalid si
                fin = true;
                if (sc.env >= 0 && !sc.hide_emitters) Lr = thr * emitter_eval(sc, sc.env, -L.ray.d.z, L.ray.d);
            } else { keep = true; working = false; }
        }
        const V2 fp = box ? L.pos : L.sample_pos;
        wv_finish<C, TO_FILM>(sc, a, fin, make_float4(__uint_as_float((uint32_t) lane), __uint_as_float((uint32_t) (lane >> 32)), fp.x, fp.y), Lr, f);
        if (fin) { ++pass; if (pass < a.n_passes) L.smp.advance(a.spp_per_pass); else working = false; }
    }
    if (!own_slot) {                                                                          // one atomic per warp
        const unsigned m = __ballot_sync(0xffffffffu, keep), l = threadIdx.x & 31u; uint32_t base = 0;
        if (m) {
            if (l == (unsigned) (__ffs(m) - 1)) base = atomicAdd(&a.q->n_alloc, (uint32_t) __popc(m));
            base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
            slot = base + __popc(m & ((1u << l) - 1u));
        }
    }
    if (keep) {
        const WVPlanes &p = a.p; const V2 fp = box ? L.pos : L.sample_pos;
        wv_rng_store(p, slot, L.smp);
        p.o(slot) = make_float4(L.ray.o.x, L.ray.o.y, L.ray.o.z, L.ray.maxt);
        p.d(slot) = make_float4(L.ray.d.x, L.ray.d.y, L.ray.d.z, __uint_as_float(f));
        p.thr(slot) = make_float4(thr.x, thr.y, thr.z, __uint_as_float(0u));
        p.hit(slot) = wf_hit4(h);
        p.res(slot) = make_float4(0.f, 0.f, 0.f, 1.f);
        p.lsp(slot) = make_float4(0.f, 0.f, 0.f, 1.f);
        p.lane(slot) = make_float4(__uint_as_float((uint32_t) lane), __uint_as_float((uint32_t) (lane >> 32)), fp.x, fp.y);
        if (sc.integ.kind == LR_INTEGRATOR_BIOVOLPATH) p.adj(slot) = make_float4(LR_INF, 0.f, 0.f, 0.f);   // { previous si.t, tissue depth } (lr_wavefront_bio.cuh)
        if (sc.integ.kind == LR_INTEGRATOR_VOLPATHMIS) {                                                     // p_over_f = p_over_f_nee = 1 (lr_wavefront_mis.cuh)
            float4 *m = a.mis + (size_t) slot * 8u;
#pragma unroll
            for (int k = 0; k < 6; ++k) m[k] = make_float4(1.f, 1.f, 1.f, 0.f);
        }
    }
    wv_route(sc, a, slot, keep, f, h);
}

template <int C, bool TO_FILM>
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wv_gen(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0; unsigned long long nsamples = 0;
    const unsigned l = threadIdx.x & 31u; const uint64_t n_total = a.lane_end - a.lane_begin;
    for (;;) {
        unsigned long long base = ~0ull;
        if (l == 0) {
            const uint32_t used = *(volatile uint32_t *) &a.q->n_alloc;
            if (used + a.slack <= a.cap && *(volatile unsigned long long *) a.lane_next < n_total) base = atomicAdd(a.lane_next, 32ull);
        }
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base == ~0ull || base >= n_total) break;
        const bool have = base + l < n_total; const uint64_t lane = lane_global(a.map, a.lane_begin + base + l);
        Sampler smp; smp.state = smp.inc = 0;
        if (have) seed_sampler(sc, smp, a.seed, lane, a.spp_per_pass, a.n_passes);
        wv_start_paths<C, TO_FILM>(sc, a, have, lane, 0u, smp, false, 0u, cnt, nsamples);
    }
    flush_counters(cnt, nsamples, a.stats);
}

// later passes of lanes whose path has ended (multi-pass renders only)
template <int C, bool TO_FILM>
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wv_regen(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0; unsigned long long nsamples = 0;
    WF_FOR_QUEUE(a.q->n_regen) {
        const bool item = qi < _n; const uint32_t slot = item ? a.q_regen[qi] : 0;
        Sampler smp; smp.state = smp.inc = 0; uint64_t lane = 0; uint32_t pass = 0;
        if (item) {
            smp = wv_rng(a.p, slot); smp.advance(a.spp_per_pass);
            const float4 lp = a.p.lane(slot); lane = ((uint64_t) __float_as_uint(lp.y) << 32) | __float_as_uint(lp.x);
            pass = (__float_as_uint(a.p.d(slot).w) >> WF_PASS_SHIFT) + 1u;
        }
        wv_start_paths<C, TO_FILM>(sc, a, item, lane, pass, smp, true, slot, cnt, nsamples);
    }
    flush_counters(cnt, nsamples, a.stats);
}

// ------------------------------------------------------------------------------------------
// TRACE: persistent ray queries over a queue of slots (lr_trace.cuh).
//   MODE 0  closest hit of the path's next ray, then routing (wv_route)
//   MODE 1  shadow ray, any hit decides (production tracking, no index-matched boundaries in the scene)
//   MODE 2  shadow ray, closest hit kept for the NEE walk (reference-order tracking / scenes with null boundaries)
// ------------------------------------------------------------------------------------------
struct TraceCfg { uint32_t n_smem, sd; int *ovf; uint32_t ovf_stride; };   // nodes held in shared memory, shared stack depth, overflow scratch

template <int MODE, typename STK>
__global__ void __launch_bounds__(LR_TRACE_THREADS, 1) k_wv_trace(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a, const __grid_constant__ TraceCfg cfg) {
    // dynamic shared memory: [staged next rays: 2 x T float4][nodes: n_smem x 4 uint4, chunk-rotated][stack: sd x T STK]
    extern __shared__ __align__(16) unsigned char s_raw[];
    float4 *s_next = reinterpret_cast<float4 *>(s_raw);
    uint4 *s_nodes = reinterpret_cast<uint4 *>(s_raw + (size_t) 2 * LR_TRACE_THREADS * sizeof(float4));
    TraceStack<STK> st; st.stack = reinterpret_cast<STK *>(s_nodes + (size_t) cfg.n_smem * 4u) + threadIdx.x; st.stride = LR_TRACE_THREADS; st.sd = (int) cfg.sd;
    st.ovf = cfg.ovf + (size_t) blockIdx.x * LR_TRACE_THREADS + threadIdx.x; st.ovf_stride = (int) cfg.ovf_stride;
    for (uint32_t i = threadIdx.x; i < cfg.n_smem * 4u; i += LR_TRACE_THREADS) {          // chunk k of node n goes to slot (k + (n >> 1)) & 3
        const uint32_t n = i >> 2, k = i & 3u;
        s_nodes[n * 4u + ((k + (n >> 1)) & 3u)] = reinterpret_cast<const uint4 *>(sc.nodes4)[i];
    }
    __syncthreads();
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    const uint32_t *q_in = MODE == 0 ? a.q_trace : a.q_shadow;
    const uint32_t n_items = MODE == 0 ? a.q->n_trace : a.q->n_shadow;
    uint32_t *cursor = &a.q->cursor[MODE == 0 ? 0 : 2];
    const unsigned lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    TraceRay r; r.node = TR_IDLE; r.sp = 0; r.tmax = 0.f; r.o = r.d = r.inv = r.ood = mk3(0.f); r.h.t = LR_INF; r.h.prim = 0xffffffffu; r.h.u = r.h.v = 0.f;
    uint32_t slot = 0; bool exhausted = false;
    // queue entries are handed out in chunks per warp: one global atomic per chunk instead of one per refill
    const uint32_t n_warps = gridDim.x * (blockDim.x >> 5);
    const uint32_t chunk = min(256u, max(32u, (n_items / (4u * n_warps) + 31u) & ~31u));
    uint32_t chunk_next = 0, chunk_end = 0, slot_a = 0, slot_b = 0, flags = 0; bool have_a = false, have_b = false;
    for (;;) {
        const unsigned walking = __ballot_sync(0xffffffffu, trace_walking(r));
        if (__popc(walking) <= 32 - LR_TRACE_REFILL || (exhausted && __any_sync(0xffffffffu, r.node == TR_DONE || (r.node == TR_IDLE && (have_a || have_b))))) {
            // ---- parked lanes: write the result out, hand the path to its next stage ----
            const bool done = r.node == TR_DONE;
            if (MODE == 0) {
                if (done) a.p.hit(slot) = wf_hit4(r.h);
                wv_route(sc, a, slot, done, flags, r.h);
            } else {
                bool walk = false;
                if (done) {
                    bool blocked = false;
                    if (MODE == 1) blocked = hit_valid(r.h);
                    else if (sc.integ.tracking == LR_TRACKING_DDA && hit_valid(r.h)) {
                        // production tracking: a non-null surface anywhere on the segment makes the transmittance exactly 0
                        // (eval_null_transmission, volpath.cpp:528-543), so the walk is skipped; only the RNG consumption differs
                        blocked = true;
                        if (sc.any_null_bsdf) {
                            const float4 o4 = a.p.no(slot), d4 = a.p.nd(slot); Ray ray; ray.o = mk3(o4.x, o4.y, o4.z); ray.d = mk3(d4.x, d4.y, d4.z); ray.maxt = o4.w;
                            cnt.rays++; blocked = hit_valid(bvh_trace<true, true>(sc, ray));
                        }
                    }
                    if (!blocked) {
                        const float4 w = a.p.nw(slot);
                        if (__float_as_int(w.w) < 0 && !hit_valid(r.h)) {      // nothing but vacuum up to the emitter: transmittance 1
                            float4 r4 = a.p.res(slot); r4.x += w.x; r4.y += w.y; r4.z += w.z; a.p.res(slot) = r4;
                        } else { walk = true; if (MODE == 2) a.p.nhit(slot) = wf_hit4(r.h); }
                    }
                }
                wf_push(a.q_nee, &a.q->n_nee, slot, walk);
            }
            if (done) r.node = TR_IDLE;
            // ---- refill, a three-stage pipeline per lane so that no stage waits for DRAM:
            //      (c) an idle lane starts the ray that was staged in shared memory by an earlier refill,
            //      (b) a lane whose staging buffer is free starts the asynchronous copy (cp.async) of the ray of the queue entry it holds,
            //      (a) a lane that holds no queue entry claims the next one (plain load, consumed at a later refill) ----
            if (r.node == TR_IDLE && have_b) {
                __pipeline_wait_prior(0);
                const float4 o4 = s_next[threadIdx.x], d4 = s_next[LR_TRACE_THREADS + threadIdx.x];
                slot = slot_b; have_b = false; flags = __float_as_uint(d4.w);
                trace_begin(sc, r, mk3(o4.x, o4.y, o4.z), mk3(d4.x, d4.y, d4.z), o4.w); cnt.rays++;
            }
            if (!have_b && have_a) {
                slot_b = slot_a; have_a = false; have_b = true;
                __pipeline_memcpy_async(&s_next[threadIdx.x], MODE == 0 ? &a.p.o(slot_b) : &a.p.no(slot_b), 16);
                __pipeline_memcpy_async(&s_next[LR_TRACE_THREADS + threadIdx.x], MODE == 0 ? &a.p.d(slot_b) : &a.p.nd(slot_b), 16);
                __pipeline_commit();
            }
            if (!exhausted) {
                const unsigned need = __ballot_sync(0xffffffffu, !have_a); const uint32_t want = (uint32_t) __popc(need);
                if (want) {
                    if (chunk_next == chunk_end) {
                        uint32_t base = 0;
                        if (lane == 0) base = atomicAdd(cursor, chunk);
                        base = __shfl_sync(0xffffffffu, base, 0);
                        chunk_next = base; chunk_end = base + chunk;
                    }
                    const uint32_t take = min(want, chunk_end - chunk_next);
                    const uint32_t rank = (uint32_t) __popc(need & lt), qi = chunk_next + rank;
                    if (!have_a && rank < take && qi < n_items) { slot_a = q_in[qi]; have_a = true; }
                    chunk_next += take;
                    exhausted = chunk_next >= n_items;
                }
            }
        }
        // ---- one step: either the lanes holding a triangle test it, or the lanes at an inner node visit it ----
        const bool at_leaf = r.node < 0;
        const unsigned leaves = __ballot_sync(0xffffffffu, at_leaf), inner = __ballot_sync(0xffffffffu, trace_walking(r) && !at_leaf);
        if (leaves != 0u && ((uint32_t) __popc(leaves) >= LR_TRACE_LEAF_BATCH || inner == 0u)) {
            if (at_leaf) trace_leaf<MODE == 1, false, STK>(sc, r, st);
        } else if ((inner >> lane) & 1u) trace_visit<STK>(sc, r, s_nodes, cfg.n_smem, st);
        if (exhausted && !__any_sync(0xffffffffu, r.node != TR_IDLE || have_a || have_b)) break;
    }
    flush_counters(cnt, 0, a.stats);
}

// ------------------------------------------------------------------------------------------
// TRACK: volpath.cpp:160-253 for paths inside a medium.  The russian roulette that opens the iteration was drawn by the
// stage that spawned the ray; further loop iterations of the reference (after a null collision) draw theirs here.
// Events: 0 end of path, 1 real scatter, 2 boundary / escape.
// ------------------------------------------------------------------------------------------
template <int C, bool TO_FILM>
LR_DEV void wv_track_out(const DScene &sc, const WVArgs &a, uint32_t slot, int event, const Sampler &smp, const Ray &ray, V3 thr, uint32_t depth, float hit_t,
                         uint32_t f, const Hit &h0) {
    const WVPlanes &p = a.p;
    const bool fin = event == 0; float4 lp = make_float4(0.f, 0.f, 0.f, 0.f); V3 L = mk3(0.f);
    if (event >= 0) wv_rng_store(p, slot, smp);
    if (fin) { lp = p.lane(slot); const float4 r4 = p.res(slot); L = mk3(r4.x, r4.y, r4.z); }
    wv_finish<C, TO_FILM>(sc, a, fin, lp, L, f);
    if (event > 0) {
        p.o(slot) = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.maxt);
        p.thr(slot) = make_float4(thr.x, thr.y, thr.z, __uint_as_float(depth));
        if (event == 3) p.d(slot) = make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float(f));      // scatter already shaded (wv_track_scatter_sealed): new direction
        else p.hit(slot).x = hit_t;
    }
    wf_push(a.q_trace, &a.q->n_trace, slot, event == 3);
    const bool esc = event == 2 && !hit_valid(h0);
    const uint32_t b = (event == 2 && !esc) ? wv_bucket(sc, h0) : 0xffu;
    wf_push(a.q_regen, &a.q->n_regen, slot, fin && wv_more_passes(a, f));
    wf_push(a.q_scatter, &a.q->n_scatter, slot, event == 1);
    wf_push(a.q_escape, &a.q->n_escape, slot, esc);
#pragma unroll
    for (uint32_t k = 0; k < WF_N_SURF; ++k) wf_push(a.q_surf[k], &a.q->n_surf[k], slot, b == k);
}

template <int C, bool TO_FILM>
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wv_track(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    const WVPlanes &p = a.p;
    WF_FOR_QUEUE(a.q->n_track) {
        const bool item = qi < _n; const uint32_t slot = item ? a.q_track[qi] : 0;
        int event = -1; Sampler smp; smp.state = smp.inc = 0; Ray ray; ray.o = ray.d = mk3(0.f); ray.maxt = 0.f; V3 thr = mk3(0.f);
        uint32_t depth = 0, f = 0; float hit_t = 0.f; Hit h0; h0.t = LR_INF; h0.prim = 0xffffffffu; h0.u = h0.v = 0.f;
        if (item) {
            smp = wv_rng(p, slot);
            float4 t4 = p.o(slot); ray.o = mk3(t4.x, t4.y, t4.z); ray.maxt = t4.w;
            t4 = p.d(slot); ray.d = mk3(t4.x, t4.y, t4.z); f = __float_as_uint(t4.w);
            t4 = p.thr(slot); thr = mk3(t4.x, t4.y, t4.z); depth = __float_as_uint(t4.w);
            h0 = wf_hit(p.hit(slot)); hit_t = h0.t;
            const float eta = p.res(slot).w;
            const uint32_t channel = f & WF_CHANNEL_MASK; const int medium = wf_medium(f);
            const DMedium &m = sc.media[medium];
            for (bool first = true;; first = false) {
                if (!first && !wv_rr(sc, smp, thr, eta, depth)) { event = 0; break; }
                if (medium_uses_dda(sc, m)) {
                    // production tracking: delta tracking against the local majorants of the super-grid.  The majorant is one
                    // scalar for all channels (heterogeneous.cpp:163), so Tr/pdf reduces to 1/mu at a collision and 1 at a surface.
                    const DGrid &g = sc.grids[m.m.sigma_t_grid];
                    float mint, maxt, t_last = 0.f;
                    if (medium_clip(g, ray, mint, maxt)) {
                        const float t_end = fminf(maxt, hit_t); t_last = mint;
                        if (mint < t_end) {
                            DDA dd; dda_init(g, ray, mint, dd); float t = mint, tau = -logf(1.f - smp.next_1d()), mu;
                            while (dda_walk(g, m.m.scale, dd, t, tau, t_end, mu)) {
                                cnt.collisions++;
                                V3 pt = ray_at(ray, t);
                                V3 st = grid_eval<false>(g, pt, nullptr) * m.m.scale, sn = mk3(mu) - st;
                                float pn = hmean(sn / mu);
                                if (smp.next_1d() < pn) { thr *= sn / (mu * pn); tau = -logf(1.f - smp.next_1d()); t_last = t; continue; }
                                depth += 1; cnt.real_scatters++; p.lsp(slot) = make_float4(pt.x, pt.y, pt.z, p.lsp(slot).w);
                                if (!(depth < sc.max_depth)) { event = 0; break; }
                                thr *= (st * medium_albedo(sc, m, pt)) / (mu * hmean(st / mu));
                                ray.o = pt; hit_t = hit_t - t; event = 1; break;
                            }
                        }
                    }
                    if (event < 0) { event = 2; if (reference_step_underflows(m, t_last, hit_t)) thr = mk3(0.f); }
                    break;
                }
                MediumInteraction mei = sample_interaction<false>(sc, medium, ray, smp.next_1d(), channel, nullptr, cnt.collisions);
                if (hit_t < mei.t) mei.t = LR_INF;
                const bool is_spectral = m.m.has_spectral_extinction != 0;
                if (is_spectral) {
                    float t = fminf(mei.t, hit_t) - mei.mint;
                    V3 tr = exp3(mei.combined_extinction * (-t));
                    V3 pdf = hit_t < mei.t ? tr : tr * mei.combined_extinction;
                    float tr_pdf = idx3(pdf, channel);
                    thr *= tr_pdf > 0.f ? tr / tr_pdf : mk3(0.f);
                }
                if (mei_valid(mei)) {
                    float null_scatter_prob = hmean(mei.sigma_n / mei.combined_extinction);
                    if (smp.next_1d() < null_scatter_prob) {                       // null collision: next loop iteration of the reference
                        if (is_spectral) thr *= mei.sigma_n / null_scatter_prob;
                        ray.o = mei.p; hit_t = hit_t - mei.t;
                        continue;
                    }
                    depth += 1; cnt.real_scatters++; p.lsp(slot) = make_float4(mei.p.x, mei.p.y, mei.p.z, p.lsp(slot).w);
                    if (!(depth < sc.max_depth)) { event = 0; break; }
                    if (is_spectral) thr *= mei.sigma_s / hmean(mei.sigma_t / mei.combined_extinction);
                    else thr *= mei.sigma_s / mei.sigma_t;
                    ray.o = mei.p; event = 1; break;
                }
                event = 2; break;
            }
        }
        wv_track_out<C, TO_FILM>(sc, a, slot, event, smp, ray, thr, depth, hit_t, f, h0);
    }
    flush_counters(cnt, 0, a.stats);
}

// A real scatter inside a SEALED medium in production tracking (DMedium::sealed: every shadow ray from inside ends on an opaque surface) needs no
// SHADE_A -- the emitter sample is two discarded draws (wv_nee_setup) -- so TRACK finishes the vertex itself: flags, phase sample, the roulette of
// the next iteration (k_wv_shade_a_medium + k_wv_shade_b_medium, same draws in the same order), and hands the path straight to TRACE_C (event 3)
// instead of sending its state through two more kernels (10.8 % of the C3 step).  Not inlined: the walk loop keeps its register budget.
__noinline__ LR_DEV void wv_track_scatter_sealed(const DScene &sc, const WVPlanes &p, uint32_t slot, Sampler &smp, Ray &ray, V3 &thr, uint32_t depth, uint32_t &f, int &event) {
    const DMedium &m = sc.media[wf_medium(f)];
    if (!m.sealed) return;
    const bool sample_emitters = m.m.sample_emitters != 0;
    f |= WF_VALID_RAY; f = sample_emitters ? (f & ~WF_SPEC_CHAIN) : (f | WF_SPEC_CHAIN);
    if (sample_emitters) (void) smp.next_2d();
    Phase phase; phase.kind = m.m.phase; phase.g = m.m.phase_g;
    (void) smp.next_1d(); const V2 s2 = smp.next_2d();
    float phase_pdf; const V3 wo = phase.sample(-ray.d, s2, phase_pdf);
    const float eta = depth > (uint32_t) sc.rr_depth ? p.res(slot).w : 1.f;
    if (wv_rr(sc, smp, thr, eta, depth)) {
        if (phase_pdf > 0.f) { ray.d = wo; ray.maxt = LR_LARGEST; reinterpret_cast<float *>(&p.lsp(slot))[3] = phase_pdf; }
        event = 3;
    } else event = 0;
}

// TRACK for scenes whose media all use the majorant DDA (production mode): the same per-item work as k_wv_track, run as a
// persistent "while-while" loop.  The number of super-voxels a free flight crosses is roughly exponentially distributed, so a
// warp that waits for its longest walk keeps 4-5 of 32 lanes busy; here a lane that has finished its item parks, and once
// half of the warp is parked the parked lanes are written out and pull the next queue entries (one atomicAdd on the cursor
// per refill).  Per-lane RNG draw order is unchanged: first optical depth, then (null test, next optical depth) per collision.
#ifndef LR_TRACK_REFILL
#define LR_TRACK_REFILL 16            // refill when at most this many lanes of the warp are still walking
#endif
#ifndef LR_TRACK_STEPS
#define LR_TRACK_STEPS 6              // super-voxel steps per lane between two converged points
#endif
// SMAJ: the majorant super-grid of grid `smaj_grid` (`smaj_floats` values, a multiple of 4) is staged in shared memory by ONE bulk
// asynchronous copy through the TMA unit at kernel start (lr_tma.cuh) -- every DDA step then reads its majorant with an LDS instead of a
// dependent global load (the walk is a chain of such loads: the kernel is latency-bound, ncu long-scoreboard).  128 KiB for the 256^3
// grid at cell 8, so the block is 1024 threads, one per SM (same resident threads as 4 x 256).  Grids whose super-grid does not fit
// (C3-XL: 1 MiB) and scenes with several grids use the global-memory form.
struct TrackSmaj { int grid; uint32_t floats; int fuse; };      // fuse: finish scatter vertices of sealed media inside TRACK (wv_track_scatter_sealed)
template <int C, bool TO_FILM, bool SMAJ>
__global__ void __launch_bounds__(SMAJ ? 1024 : 256, SMAJ ? 1 : LR_WF_OCC) k_wv_track_dda(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a, const TrackSmaj sm) {
    extern __shared__ __align__(16) float s_maj[];
    if (SMAJ) {
        __shared__ __align__(8) uint64_t bar;
        if (threadIdx.x == 0) tma_bar_init(&bar, 1);
        __syncthreads();
        if (threadIdx.x == 0) {
            const uint32_t bytes = sm.floats * 4u; tma_bar_expect(&bar, bytes);
            const char *src = reinterpret_cast<const char *>(sc.grids[sm.grid].maj); char *dst = reinterpret_cast<char *>(s_maj);
            for (uint32_t off = 0; off < bytes; off += 32768u) tma_load_1d(dst + off, src + off, min(32768u, bytes - off), &bar);
        }
        tma_bar_wait(&bar, 0);
    }
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    const WVPlanes &p = a.p;
    const uint32_t n_items = a.q->n_track;
    const unsigned lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    uint32_t slot = 0, depth = 0, f = 0; int event = -2;                  // -2 idle, -1 walking, 0 end of path, 1 scatter, 2 boundary
    Sampler smp; smp.state = smp.inc = 0; Ray ray; ray.o = ray.d = mk3(0.f); ray.maxt = 0.f; V3 thr = mk3(0.f);
    float hit_t = 0.f, t = 0.f, tau = 0.f, t_end = 0.f, mu_scale = 0.f, t_last = 0.f; int grid_id = 0; DDA dd{};
    Hit h0; h0.t = LR_INF; h0.prim = 0xffffffffu; h0.u = h0.v = 0.f;
    bool exhausted = false;
    for (;;) {
        const unsigned walking = __ballot_sync(0xffffffffu, event == -1);
        if (__popc(walking) <= LR_TRACK_REFILL) {
            // ---- write finished items out and queue them (converged: the pushes use warp ballots) ----
            if (__any_sync(0xffffffffu, event >= 0)) {
                if (event == 1 && sm.fuse) wv_track_scatter_sealed(sc, p, slot, smp, ray, thr, depth, f, event);
                wv_track_out<C, TO_FILM>(sc, a, slot, event >= 0 ? event : -1, smp, ray, thr, depth, hit_t, f, h0);
                if (event >= 0) event = -2;
            }
            if (!exhausted) {
                const unsigned idle = ~walking; uint32_t base = 0;
                if (lane == (unsigned) (__ffs(idle) - 1)) base = atomicAdd(&a.q->cursor[1], (uint32_t) __popc(idle));
                base = __shfl_sync(0xffffffffu, base, __ffs(idle) - 1);
                exhausted = base + (uint32_t) __popc(idle) >= n_items;
                const uint32_t qi = base + (uint32_t) __popc(idle & lt);
                if (event == -2 && qi < n_items) {
                    slot = a.q_track[qi];
                    smp = wv_rng(p, slot);
                    float4 t4 = p.o(slot); ray.o = mk3(t4.x, t4.y, t4.z); ray.maxt = t4.w;
                    t4 = p.d(slot); ray.d = mk3(t4.x, t4.y, t4.z); f = __float_as_uint(t4.w);
                    t4 = p.thr(slot); thr = mk3(t4.x, t4.y, t4.z); depth = __float_as_uint(t4.w);
                    h0 = wf_hit(p.hit(slot)); hit_t = h0.t;
                    const DMedium &m = sc.media[wf_medium(f)]; grid_id = m.m.sigma_t_grid; mu_scale = m.m.scale; const DGrid &g = sc.grids[grid_id];
                    event = 2; t_last = 0.f;
                    float mint, maxt;
                    if (medium_clip(g, ray, mint, maxt)) {
                        t_end = fminf(maxt, hit_t); t_last = mint;
                        if (mint < t_end) { dda_init(g, ray, mint, dd); t = mint; tau = -logf(1.f - smp.next_1d()); event = -1; }
                    }
                    if (event == 2 && reference_step_underflows(m, t_last, hit_t)) thr = mk3(0.f);
                }
            }
        }
        // ---- walk: bring every walking lane to its next tentative collision (or the end of its segment) ----
        bool collide = false; float mu = 0.f;
        if (event == -1) {
            const DGrid &g = sc.grids[grid_id];
#pragma unroll 1
            for (int k = 0; k < LR_TRACK_STEPS; ++k) {
                mu = mu_scale * ((SMAJ && grid_id == sm.grid) ? s_maj[((size_t) dd.iz * g.my + dd.iy) * g.mx + dd.ix] : dda_majorant(g, dd));
                const float t_exit = fminf(fminf(dd.tnx, dd.tny), fminf(dd.tnz, t_end));
                const float dtau = mu * (t_exit - t);
                if (dtau >= tau && mu > 0.f) { t = fminf(t + tau / mu, t_exit); collide = true; break; }
                tau -= dtau; t = t_exit;
                if (t_exit >= t_end) { event = 2; if (reference_step_underflows(sc.media[wf_medium(f)], t_last, hit_t)) thr = mk3(0.f); break; }
                dda_advance(g, dd);
            }
        }
        // ---- tentative collisions, evaluated together ----
        if (collide) {
            const DGrid &g = sc.grids[grid_id]; const DMedium &m = sc.media[wf_medium(f)];
            cnt.collisions++;
            V3 pt = ray_at(ray, t);
            V3 st = grid_eval<false>(g, pt, nullptr) * mu_scale, sn = mk3(mu) - st;
            float pn = hmean(sn / mu);
            if (smp.next_1d() < pn) { thr *= sn / (mu * pn); tau = -logf(1.f - smp.next_1d()); t_last = t; }
            else {
                depth += 1; cnt.real_scatters++;
                float *lp = reinterpret_cast<float *>(&p.lsp(slot)); lp[0] = pt.x; lp[1] = pt.y; lp[2] = pt.z;
                if (!(depth < sc.max_depth)) event = 0;
                else { thr *= (st * medium_albedo(sc, m, pt)) / (mu * hmean(st / mu)); ray.o = pt; hit_t = hit_t - t; event = 1; }
            }
        }
        if (exhausted && !__any_sync(0xffffffffu, event != -2)) break;
    }
    flush_counters(cnt, 0, a.stats);
}

// ------------------------------------------------------------------------------------------
// SHADE_A: emitter sampling (Scene::sample_emitter_direction + the first lines of VolumetricPathIntegrator::sample_emitter,
// volpath.cpp:400-452).  The shadow ray is written to the NEE planes; TRACE_S / NEE finish the estimate.
// ------------------------------------------------------------------------------------------
LR_DEV bool wv_nee_setup(const DScene &sc, const WVArgs &a, uint32_t slot, Sampler &smp, V3 ref_p, V3 ref_n, const SurfaceInteraction *ref_si,
                         int &medium, DirectionSample &ds, V3 &emitter_val) {
    // production tracking: a shadow ray that starts in a sealed medium always ends on an opaque surface (DMedium::sealed) => exactly 0;
    // the two draws of the emitter sample are consumed as usual, nothing else happens
    const bool dda = sc.integ.tracking == LR_TRACKING_DDA;
    if (dda && !ref_si && medium >= 0 && sc.media[medium].sealed) { (void) smp.next_2d(); ds.pdf = 0.f; return false; }
    ds = sample_emitter_direction(sc, ref_p, smp.next_2d(), emitter_val);
    if (ds.pdf == 0.f) return false;
    const Ray ray = spawn_ray_to(ref_p, ref_n, ds.p);
    if (ref_si && mesh_is_transition(sc, ref_si->mesh)) medium = target_medium(sc, *ref_si, ray.d);
    if (dda && medium >= 0 && sc.media[medium].sealed) return false;
    a.p.no(slot) = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.maxt);
    a.p.nd(slot) = make_float4(ray.d.x, ray.d.y, ray.d.z, ds.dist);
    return true;
}

// medium scatter (volpath.cpp:256-283)
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wv_shade_a_medium(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a) {
    const WVPlanes &p = a.p;
    WF_FOR_QUEUE(a.q->n_scatter) {
        const bool item = qi < _n; bool nee = false; const uint32_t slot = item ? a.q_scatter[qi] : 0;
        if (item) {
            const float4 d4 = p.d(slot); uint32_t f = __float_as_uint(d4.w); const int medium = wf_medium(f);
            const DMedium &m = sc.media[medium]; Phase phase; phase.kind = m.m.phase; phase.g = m.m.phase_g;
            const bool sample_emitters = m.m.sample_emitters != 0;
            const uint32_t f0 = f; int walk_medium = medium;
            f |= WF_VALID_RAY; f = sample_emitters ? (f & ~WF_SPEC_CHAIN) : (f | WF_SPEC_CHAIN);
            if (sample_emitters) {
                Sampler smp = wv_rng(p, slot);
                const float4 o4 = p.o(slot), t4 = p.thr(slot); const V3 pt = mk3(o4.x, o4.y, o4.z), wi = -mk3(d4.x, d4.y, d4.z), thr = mk3(t4.x, t4.y, t4.z);
                DirectionSample ds; V3 ev;
                if (wv_nee_setup(sc, a, slot, smp, pt, mk3(0.f), nullptr, walk_medium, ds, ev)) {
                    const float phase_val = phase.eval_pdf(wi, ds.d);
                    const V3 w = thr * phase_val * mis_weight(ds.pdf, ds.delta ? 0.f : phase_val) * ev;
                    p.nw(slot) = make_float4(w.x, w.y, w.z, __int_as_float(walk_medium));
                    nee = wv_finite(w);
                }
                wv_rng_store(p, slot, smp);
            }
            if (f != f0) reinterpret_cast<float *>(&p.d(slot))[3] = __uint_as_float(f);
        }
        wf_push(a.q_shadow, &a.q->n_shadow, slot, nee);
    }
}

// surface (volpath.cpp:300-362); every item has a valid hit (escaped rays have their own queue)
template <int C, bool TO_FILM>
__global__ void __launch_bounds__(128, LR_WF_OCC_SURF) k_wv_shade_a_surface(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a, int bucket) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    const WVPlanes &p = a.p;
    WF_FOR_QUEUE(a.q->n_surf[bucket]) {
        const bool item = qi < _n; const uint32_t slot = item ? a.q_surf[bucket][qi] : 0; bool nee = false, gone = false; uint32_t f = 0;
        if (item) {
            Sampler smp = wv_rng(p, slot);
            const float4 o4 = p.o(slot), d4 = p.d(slot), t4 = p.thr(slot);
            f = __float_as_uint(d4.w); const uint32_t depth = __float_as_uint(t4.w); const int medium = wf_medium(f);
            Ray ray; ray.o = mk3(o4.x, o4.y, o4.z); ray.d = mk3(d4.x, d4.y, d4.z); ray.maxt = o4.w;
            const V3 thr = mk3(t4.x, t4.y, t4.z);
            Hit h = wf_hit(p.hit(slot));
            SurfaceInteraction si = compute_si(sc, ray, h, sc.tris_gid);
            // the segment started outside any medium <=> the reference intersected in this very iteration (`intersect`, :301-303)
            if (sc.hide_emitters && medium < 0 && depth == 0u && sc.meshes[si.mesh].emitter >= 0) {                     // :307-323
                Ray r2 = spawn_ray(si.p, si.n, ray.d); Hit pi = skip_area_emitters(sc, sc.tris_gid, r2, cnt);
                si = compute_si(sc, r2, pi, sc.tris_gid); h = pi; p.hit(slot) = wf_hit4(h);
                p.o(slot) = make_float4(r2.o.x, r2.o.y, r2.o.z, r2.maxt);         // SHADE_B recomputes si from (ray, hit): keep them consistent
            }
            const bool count_direct = depth == 0u || (f & WF_SPEC_CHAIN);
            const int emitter = emitter_of(sc, si);
            if (emitter >= 0 && !(depth == 0u && sc.hide_emitters)) {                                                     // :326-345
                const float4 l4 = p.lsp(slot);
                float emitter_pdf = 1.f;
                if (!count_direct) { DirectionSample ds = ds_from_si(sc, si, mk3(l4.x, l4.y, l4.z)); emitter_pdf = pdf_emitter_direction(sc, ds); }
                const V3 emitted = eval_emitter_at(sc, emitter, si);
                const V3 c = count_direct ? thr * emitted : thr * mis_weight(l4.w, emitter_pdf) * emitted;
                if (wv_finite(c)) { float4 r4 = p.res(slot); r4.x += c.x; r4.y += c.y; r4.z += c.z; p.res(slot) = r4; }
            }
            if (!si_valid(si)) {                                   // only after the hidden-emitter skip above: nothing behind the emitter
                gone = true; f |= WF_DEAD; reinterpret_cast<float *>(&p.d(slot))[3] = __uint_as_float(f);
            } else {
                BSDF bsdf(sc, si);
                if ((bsdf.flags() & F_SMOOTH) && (depth + 1 < sc.max_depth)) {                                           // :350-362
                    DirectionSample ds; V3 ev; int walk_medium = medium;
                    if (wv_nee_setup(sc, a, slot, smp, si.p, si.n, &si, walk_medium, ds, ev)) {
                        const V3 wo = si.sh.to_local(ds.d); float bsdf_pdf; const V3 bsdf_val = bsdf.eval_pdf(si.wi, wo, bsdf_pdf);
                        const V3 w = thr * bsdf_val * mis_weight(ds.pdf, ds.delta ? 0.f : bsdf_pdf) * ev;
                        p.nw(slot) = make_float4(w.x, w.y, w.z, __int_as_float(walk_medium));
                        nee = wv_finite(w);
                    }
                    wv_rng_store(p, slot, smp);
                }
            }
        }
        {
            float4 lp = make_float4(0.f, 0.f, 0.f, 0.f); V3 L = mk3(0.f);
            if (gone) { lp = p.lane(slot); const float4 r4 = p.res(slot); L = mk3(r4.x, r4.y, r4.z); }
            wv_finish<C, TO_FILM>(sc, a, gone, lp, L, f);
        }
        wf_push(a.q_regen, &a.q->n_regen, slot, gone && wv_more_passes(a, f));
        wf_push(a.q_shadow, &a.q->n_shadow, slot, nee);
    }
    flush_counters(cnt, 0, a.stats);
}

// rays that left the scene (volpath.cpp:326-345 with an invalid si): environment radiance + MIS, then the path ends
template <int C, bool TO_FILM>
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wv_escape(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a) {
    const WVPlanes &p = a.p;
    WF_FOR_QUEUE(a.q->n_escape) {
        const bool item = qi < _n; const uint32_t slot = item ? a.q_escape[qi] : 0;
        V3 L = mk3(0.f); uint32_t f = 0; float4 lp = make_float4(0.f, 0.f, 0.f, 0.f);
        if (item) {
            const float4 d4 = p.d(slot), t4 = p.thr(slot), r4 = p.res(slot); lp = p.lane(slot);
            f = __float_as_uint(d4.w); const uint32_t depth = __float_as_uint(t4.w);
            L = mk3(r4.x, r4.y, r4.z);
            if (sc.env >= 0 && !(depth == 0u && sc.hide_emitters)) {
                const V3 dir = mk3(d4.x, d4.y, d4.z), thr = mk3(t4.x, t4.y, t4.z);
                const bool count_direct = depth == 0u || (f & WF_SPEC_CHAIN);
                const V3 emitted = emitter_eval(sc, sc.env, -dir.z, dir);
                if (count_direct) L += thr * emitted;
                else {
                    DirectionSample ds; ds.d = dir; ds.emitter = sc.env; ds.dist = 0.f; ds.n = ds.p = mk3(0.f); ds.uv = mk2(0.f, 0.f); ds.pdf = 0.f; ds.delta = false;
                    L += thr * mis_weight(p.lsp(slot).w, pdf_emitter_direction(sc, ds)) * emitted;
                }
            }
        }
        wv_finish<C, TO_FILM>(sc, a, item, lp, L, f);
        wf_push(a.q_regen, &a.q->n_regen, slot, item && wv_more_passes(a, f));
    }
}

// ------------------------------------------------------------------------------------------
// NEE: transmittance along the shadow segment (volpath.cpp:454-551), ratio tracking through media and index-matched
// (null) boundaries, over the slots TRACE_S let through.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wv_nee(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a, int have_nhit) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    const WVPlanes &p = a.p;
    WF_FOR_QUEUE(a.q->n_nee) {
        if (qi < _n) {
            const uint32_t slot = a.q_nee[qi];
            Sampler smp = wv_rng(p, slot);
            const uint32_t channel = __float_as_uint(p.d(slot).w) & WF_CHANNEL_MASK;
            const float4 o4 = p.no(slot), d4 = p.nd(slot), w4 = p.nw(slot);
            Ray ray; ray.o = mk3(o4.x, o4.y, o4.z); ray.d = mk3(d4.x, d4.y, d4.z);
            const float max_dist = o4.w, ds_dist = d4.w; float total_dist = 0.f;
            int medium = __float_as_int(w4.w);
            Hit h; h.t = LR_INF; h.prim = 0xffffffffu; h.u = h.v = 0.f;
            if (have_nhit) h = wf_hit(p.nhit(slot));
            float si_t = h.t; bool needs_intersection = false, active = true;
            V3 transmittance = mk3(1.f);
            while (active) {
                float remaining_dist = max_dist - total_dist; ray.maxt = remaining_dist;
                active &= remaining_dist > 0.f;
                if (!active) break;
                bool escaped_medium = false, active_medium = medium >= 0, active_surface = !active_medium;
                if (active_medium && medium_uses_dda(sc, sc.media[medium])) {
                    // ratio tracking over the whole medium segment against the local majorants: T *= sigma_n / mu per tentative collision
                    const DMedium &m = sc.media[medium]; const DGrid &g = sc.grids[m.m.sigma_t_grid];
                    if (needs_intersection) { cnt.rays++; h = bvh_trace<false, false>(sc, ray); si_t = h.t; needs_intersection = false; }
                    float mint, maxt, t_last = 0.f;
                    if (medium_clip(g, ray, mint, maxt)) {
                        const float t_end = fminf(maxt, si_t); t_last = mint;
                        if (mint < t_end) {
                            DDA dd; dda_init(g, ray, mint, dd); float t = mint, tau = -logf(1.f - smp.next_1d()), mu;
                            while (dda_walk(g, m.m.scale, dd, t, tau, t_end, mu)) {
                                cnt.collisions++;
                                V3 st = grid_eval<false>(g, ray_at(ray, t), nullptr) * m.m.scale;
                                transmittance *= (mk3(mu) - st) / mu; t_last = t;
                                if (!any_nonzero(transmittance)) break;
                                tau = -logf(1.f - smp.next_1d());
                            }
                        }
                    }
                    if (reference_step_underflows(m, t_last, fminf(remaining_dist, si_t))) transmittance = mk3(0.f);
                    if (si_t > remaining_dist) total_dist = ds_dist;        // nothing but medium up to the emitter (volpath.cpp:494-496)
                    escaped_medium = true; active_medium = false;
                } else if (active_medium) {
                    const DMedium &m = sc.media[medium];
                    MediumInteraction mei = sample_interaction<false>(sc, medium, ray, smp.next_1d(), channel, nullptr, cnt.collisions);
                    if (needs_intersection) {
                        if (m.m.kind == LR_MEDIUM_HOMOGENEOUS && mei_valid(mei)) ray.maxt = fminf(mei.t, remaining_dist);
                        cnt.rays++; h = bvh_trace<false, false>(sc, ray); si_t = h.t; needs_intersection = false;
                    }
                    if (si_t < mei.t) mei.t = LR_INF;
                    const bool is_spectral = m.m.has_spectral_extinction != 0;
                    if (is_spectral) {
                        float t = fminf(remaining_dist, fminf(mei.t, si_t)) - mei.mint;
                        V3 tr = exp3(mei.combined_extinction * (-t));
                        V3 ffp = (si_t < mei.t || mei.t > remaining_dist) ? tr : tr * mei.combined_extinction;
                        float tr_pdf = idx3(ffp, channel);
                        transmittance *= tr_pdf > 0.f ? tr / tr_pdf : mk3(0.f);
                    }
                    if (mei.t > remaining_dist && mei_valid(mei)) total_dist = ds_dist;
                    if (mei.t > remaining_dist) mei.t = LR_INF;
                    escaped_medium = !mei_valid(mei); active_medium = mei_valid(mei);
                    if (active_medium) {
                        total_dist += mei.t; ray.o = mei.p; si_t = si_t - mei.t;
                        if (is_spectral) transmittance *= mei.sigma_n; else transmittance *= mei.sigma_n / mei.combined_extinction;
                    }
                }
                if (active_surface && needs_intersection) { cnt.rays++; h = bvh_trace<false, false>(sc, ray); si_t = h.t; needs_intersection = false; }
                active_surface |= escaped_medium;
                if (active_surface) total_dist += si_t;
                active_surface &= (si_t != LR_INF) && !active_medium;
                SurfaceInteraction si;
                if (active_surface) {
                    h.t = si_t; si = compute_si(sc, ray, h, sc.tris_gid);
                    BSDF bsdf(sc, si);
                    transmittance *= bsdf.eval_null_transmission();
                    Ray r2 = spawn_ray(si.p, si.n, ray.d); ray.o = r2.o;
                }
                ray.maxt = remaining_dist;
                needs_intersection |= active_surface;
                active &= (active_medium || active_surface) && any_nonzero(transmittance);
                if (active_surface && mesh_is_transition(sc, si.mesh)) medium = target_medium(sc, si, ray.d);
            }
            // result += throughput * (phase | bsdf) value * mis * emitter_val * transmittance     (:279-283, :356-361)
            float4 r4 = p.res(slot);
            r4.x = fmaf(w4.x, transmittance.x, r4.x); r4.y = fmaf(w4.y, transmittance.y, r4.y); r4.z = fmaf(w4.z, transmittance.z, r4.z);
            p.res(slot) = r4;
            wv_rng_store(p, slot, smp);
        }
    }
    flush_counters(cnt, 0, a.stats);
}

// ------------------------------------------------------------------------------------------
// SHADE_B: direction sampling, then the russian roulette + depth test of the next loop iteration.
// ------------------------------------------------------------------------------------------
// medium scatter: phase-function sampling (volpath.cpp:285-297)
template <int C, bool TO_FILM>
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wv_shade_b_medium(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a) {
    const WVPlanes &p = a.p;
    uint32_t *q_out = a.q_trace, *n_out = &a.q->n_trace;
    WF_FOR_QUEUE(a.q->n_scatter) {
        const bool item = qi < _n; const uint32_t slot = item ? a.q_scatter[qi] : 0;
        bool alive = false; uint32_t f = 0; V3 L = mk3(0.f); float4 lp = make_float4(0.f, 0.f, 0.f, 0.f);
        if (item) {
            Sampler smp = wv_rng(p, slot);
            const float4 d4 = p.d(slot), t4 = p.thr(slot); f = __float_as_uint(d4.w);
            const DMedium &m = sc.media[wf_medium(f)]; Phase phase; phase.kind = m.m.phase; phase.g = m.m.phase_g;
            (void) smp.next_1d(); const V2 s2 = smp.next_2d();
            float phase_pdf; const V3 wo = phase.sample(-mk3(d4.x, d4.y, d4.z), s2, phase_pdf);
            V3 thr = mk3(t4.x, t4.y, t4.z); const uint32_t depth = __float_as_uint(t4.w);
            // eta only enters the roulette once it is performed (depth > rr_depth): the h2 record is not touched before
            const float eta = depth > (uint32_t) sc.rr_depth ? p.res(slot).w : 1.f;
            alive = wv_rr(sc, smp, thr, eta, depth);
            wv_rng_store(p, slot, smp);
            if (alive) {
                if (phase_pdf > 0.f) {
                    p.d(slot) = make_float4(wo.x, wo.y, wo.z, d4.w);
                    reinterpret_cast<float *>(&p.o(slot))[3] = LR_LARGEST; reinterpret_cast<float *>(&p.lsp(slot))[3] = phase_pdf;
                }
                p.thr(slot) = make_float4(thr.x, thr.y, thr.z, t4.w);
            } else { const float4 r4 = p.res(slot); L = mk3(r4.x, r4.y, r4.z); lp = p.lane(slot); }
        }
        wv_finish<C, TO_FILM>(sc, a, item && !alive, lp, L, f);
        wf_push(q_out, n_out, slot, alive);
        wf_push(a.q_regen, &a.q->n_regen, slot, item && !alive && wv_more_passes(a, f));
    }
}

// surface: BSDF sampling, next ray, medium transition (volpath.cpp:364-389)
template <int C, bool TO_FILM>
__global__ void __launch_bounds__(128, LR_WF_OCC_SURF) k_wv_shade_b_surface(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a, int bucket) {
    const WVPlanes &p = a.p;
    uint32_t *q_out = a.q_trace, *n_out = &a.q->n_trace;
    WF_FOR_QUEUE(a.q->n_surf[bucket]) {
        bool item = qi < _n; const uint32_t slot = item ? a.q_surf[bucket][qi] : 0;
        bool alive = false; uint32_t f = 0; V3 L = mk3(0.f); float4 lp = make_float4(0.f, 0.f, 0.f, 0.f);
        float4 d4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (item) { d4 = p.d(slot); f = __float_as_uint(d4.w); item = !(f & WF_DEAD); }        // WF_DEAD: finished by SHADE_A (hidden-emitter skip)
        if (item) {
            Sampler smp = wv_rng(p, slot);
            const float4 o4 = p.o(slot), t4 = p.thr(slot), r4 = p.res(slot);
            Ray ray; ray.o = mk3(o4.x, o4.y, o4.z); ray.d = mk3(d4.x, d4.y, d4.z); ray.maxt = o4.w;
            const SurfaceInteraction si = compute_si(sc, ray, wf_hit(p.hit(slot)), sc.tris_gid);
            BSDF bsdf(sc, si);
            const float s1 = smp.next_1d(); const V2 s2 = smp.next_2d();
            V3 bsdf_val; const BSDFSample bs = bsdf.sample(si.wi, s1, s2, bsdf_val);
            V3 thr = mk3(t4.x, t4.y, t4.z) * bsdf_val; const float eta = r4.w * bs.eta; uint32_t depth = __float_as_uint(t4.w);
            ray = spawn_ray(si.p, si.n, si.sh.to_world(bs.wo));
            const bool non_null = !(bs.sampled_type & F_NULL);
            if (non_null) { depth += 1u; p.lsp(slot) = make_float4(si.p.x, si.p.y, si.p.z, bs.pdf); f |= WF_VALID_RAY; }
            bool spec = (f & WF_SPEC_CHAIN) != 0;
            spec |= non_null && (bs.sampled_type & F_DELTA); spec &= !(bs.sampled_type & F_SMOOTH);
            f = spec ? (f | WF_SPEC_CHAIN) : (f & ~WF_SPEC_CHAIN);
            if (mesh_is_transition(sc, si.mesh)) f = wf_set_medium(f, target_medium(sc, si, ray.d));
            alive = wv_rr(sc, smp, thr, eta, depth);
            wv_rng_store(p, slot, smp);
            if (alive) {
                p.o(slot) = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.maxt);
                p.d(slot) = make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float(f));
                p.thr(slot) = make_float4(thr.x, thr.y, thr.z, __uint_as_float(depth));
                if (eta != r4.w) reinterpret_cast<float *>(&p.res(slot))[3] = eta;
            } else { L = mk3(r4.x, r4.y, r4.z); lp = p.lane(slot); }
        }
        wv_finish<C, TO_FILM>(sc, a, item && !alive, lp, L, f);
        wf_push(q_out, n_out, slot, alive);
        wf_push(a.q_regen, &a.q->n_regen, slot, item && !alive && wv_more_passes(a, f));
    }
}

// ------------------------------------------------------------------------------------------
// Reference-order tracking (LR_TRACKING_REFERENCE: one global majorant, the reference's draw order) as persistent refill loops: the same
// per-item work as k_wv_track / k_wv_nee, but a path meets 25-28 tentative collisions on the C3 frame and the chains of one warp differ
// widely, so one loop turn is ONE tentative collision per walking lane, finished lanes park, and once LR_REF_REFILL lanes are parked they
// are written out and pull the next queue entries (the form k_wv_track_dda has).  Per-lane draw order is unchanged.
// ------------------------------------------------------------------------------------------
#ifndef LR_REF_REFILL
#define LR_REF_REFILL 12
#endif
template <int C, bool TO_FILM>
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wv_track_ref(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    const WVPlanes &p = a.p;
    const uint32_t n_items = a.q->n_track;
    const unsigned lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    int event = -2;                                                        // -2 idle, -1 walking, 0 end of path, 1 scatter, 2 boundary
    uint32_t slot = 0, depth = 0, f = 0; float hit_t = 0.f, eta = 1.f; bool first = true, exhausted = false;
    Sampler smp; smp.state = smp.inc = 0; Ray ray; ray.o = ray.d = mk3(0.f); ray.maxt = 0.f; V3 thr = mk3(0.f);
    Hit h0; h0.t = LR_INF; h0.prim = 0xffffffffu; h0.u = h0.v = 0.f;
    for (;;) {
        const unsigned walking = __ballot_sync(0xffffffffu, event == -1);
        if (__popc(walking) <= 32 - LR_REF_REFILL) {
            if (__any_sync(0xffffffffu, event >= 0)) {
                wv_track_out<C, TO_FILM>(sc, a, slot, event >= 0 ? event : -1, smp, ray, thr, depth, hit_t, f, h0);
                if (event >= 0) event = -2;
            }
            if (!exhausted) {
                const unsigned idle = ~walking; uint32_t base = 0;
                if (lane == (unsigned) (__ffs(idle) - 1)) base = atomicAdd(&a.q->cursor[1], (uint32_t) __popc(idle));
                base = __shfl_sync(0xffffffffu, base, __ffs(idle) - 1);
                exhausted = base + (uint32_t) __popc(idle) >= n_items;
                const uint32_t qi = base + (uint32_t) __popc(idle & lt);
                if (event == -2 && qi < n_items) {
                    slot = a.q_track[qi];
                    smp = wv_rng(p, slot);
                    float4 t4 = p.o(slot); ray.o = mk3(t4.x, t4.y, t4.z); ray.maxt = t4.w;
                    t4 = p.d(slot); ray.d = mk3(t4.x, t4.y, t4.z); f = __float_as_uint(t4.w);
                    t4 = p.thr(slot); thr = mk3(t4.x, t4.y, t4.z); depth = __float_as_uint(t4.w);
                    h0 = wf_hit(p.hit(slot)); hit_t = h0.t; eta = p.res(slot).w;
                    first = true; event = -1;
                }
            }
        }
        if (event == -1) {                                                  // one loop iteration of the reference for this lane (volpath.cpp:160-253)
            const uint32_t channel = f & WF_CHANNEL_MASK; const int medium = wf_medium(f);
            const DMedium &m = sc.media[medium];
            if (!first && !wv_rr(sc, smp, thr, eta, depth)) event = 0;
            else {
                first = false;
                MediumInteraction mei = sample_interaction<false>(sc, medium, ray, smp.next_1d(), channel, nullptr, cnt.collisions);
                if (hit_t < mei.t) mei.t = LR_INF;
                const bool is_spectral = m.m.has_spectral_extinction != 0;
                if (is_spectral) {
                    const float t = fminf(mei.t, hit_t) - mei.mint;
                    const V3 tr = exp3(mei.combined_extinction * (-t));
                    const V3 pdf = hit_t < mei.t ? tr : tr * mei.combined_extinction;
                    const float tr_pdf = idx3(pdf, channel);
                    thr *= tr_pdf > 0.f ? tr / tr_pdf : mk3(0.f);
                }
                if (!mei_valid(mei)) event = 2;
                else {
                    const float null_scatter_prob = hmean(mei.sigma_n / mei.combined_extinction);
                    if (smp.next_1d() < null_scatter_prob) {                       // null collision: next loop iteration of the reference
                        if (is_spectral) thr *= mei.sigma_n / null_scatter_prob;
                        ray.o = mei.p; hit_t = hit_t - mei.t;
                    } else {
                        depth += 1; cnt.real_scatters++; p.lsp(slot) = make_float4(mei.p.x, mei.p.y, mei.p.z, p.lsp(slot).w);
                        if (!(depth < sc.max_depth)) event = 0;
                        else {
                            if (is_spectral) thr *= mei.sigma_s / hmean(mei.sigma_t / mei.combined_extinction);
                            else thr *= mei.sigma_s / mei.sigma_t;
                            ray.o = mei.p; event = 1;
                        }
                    }
                }
            }
        }
        if (exhausted && !__any_sync(0xffffffffu, event != -2)) break;
    }
    flush_counters(cnt, 0, a.stats);
}

// the ratio-tracking walk of k_wv_nee for scenes in reference-order tracking (every item comes with TRACE_S's closest hit)
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wv_nee_ref(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    const WVPlanes &p = a.p;
    const uint32_t n_items = a.q->n_nee;
    const unsigned lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    int state = -2;                                                        // -2 idle, -1 walking, 0 walk finished
    uint32_t slot = 0, channel = 0; bool exhausted = false, needs_intersection = false;
    Sampler smp; smp.state = smp.inc = 0; Ray ray; ray.o = ray.d = mk3(0.f); ray.maxt = 0.f;
    float max_dist = 0.f, ds_dist = 0.f, total_dist = 0.f, si_t = LR_INF; int medium = -1; V3 transmittance = mk3(1.f);
    Hit h; h.t = LR_INF; h.prim = 0xffffffffu; h.u = h.v = 0.f;
    for (;;) {
        const unsigned walking = __ballot_sync(0xffffffffu, state == -1);
        if (__popc(walking) <= 32 - LR_REF_REFILL) {
            if (state == 0) {                                               // result += throughput * (phase | bsdf) value * mis * emitter_val * transmittance
                const float4 w4 = p.nw(slot); float4 r4 = p.res(slot);
                r4.x = fmaf(w4.x, transmittance.x, r4.x); r4.y = fmaf(w4.y, transmittance.y, r4.y); r4.z = fmaf(w4.z, transmittance.z, r4.z);
                p.res(slot) = r4;
                wv_rng_store(p, slot, smp);
                state = -2;
            }
            if (!exhausted) {
                const unsigned idle = ~walking; uint32_t base = 0;
                if (lane == (unsigned) (__ffs(idle) - 1)) base = atomicAdd(&a.q->cursor_nee, (uint32_t) __popc(idle));
                base = __shfl_sync(0xffffffffu, base, __ffs(idle) - 1);
                exhausted = base + (uint32_t) __popc(idle) >= n_items;
                const uint32_t qi = base + (uint32_t) __popc(idle & lt);
                if (state == -2 && qi < n_items) {
                    slot = a.q_nee[qi];
                    smp = wv_rng(p, slot);
                    channel = __float_as_uint(p.d(slot).w) & WF_CHANNEL_MASK;
                    const float4 o4 = p.no(slot), d4 = p.nd(slot);
                    ray.o = mk3(o4.x, o4.y, o4.z); ray.d = mk3(d4.x, d4.y, d4.z);
                    max_dist = o4.w; ds_dist = d4.w; total_dist = 0.f;
                    medium = __float_as_int(p.nw(slot).w);
                    h = wf_hit(p.nhit(slot)); si_t = h.t; needs_intersection = false; transmittance = mk3(1.f);
                    state = -1;
                }
            }
        }
        if (state == -1) {                                                  // one iteration of the walk (volpath.cpp:454-551)
            bool active = true;
            const float remaining_dist = max_dist - total_dist; ray.maxt = remaining_dist;
            active &= remaining_dist > 0.f;
            if (active) {
                bool escaped_medium = false, active_medium = medium >= 0, active_surface = !active_medium;
                if (active_medium) {
                    const DMedium &m = sc.media[medium];
                    MediumInteraction mei = sample_interaction<false>(sc, medium, ray, smp.next_1d(), channel, nullptr, cnt.collisions);
                    if (needs_intersection) {
                        if (m.m.kind == LR_MEDIUM_HOMOGENEOUS && mei_valid(mei)) ray.maxt = fminf(mei.t, remaining_dist);
                        cnt.rays++; h = bvh_trace<false, false>(sc, ray); si_t = h.t; needs_intersection = false;
                    }
                    if (si_t < mei.t) mei.t = LR_INF;
                    const bool is_spectral = m.m.has_spectral_extinction != 0;
                    if (is_spectral) {
                        const float t = fminf(remaining_dist, fminf(mei.t, si_t)) - mei.mint;
                        const V3 tr = exp3(mei.combined_extinction * (-t));
                        const V3 ffp = (si_t < mei.t || mei.t > remaining_dist) ? tr : tr * mei.combined_extinction;
                        const float tr_pdf = idx3(ffp, channel);
                        transmittance *= tr_pdf > 0.f ? tr / tr_pdf : mk3(0.f);
                    }
                    if (mei.t > remaining_dist && mei_valid(mei)) total_dist = ds_dist;
                    if (mei.t > remaining_dist) mei.t = LR_INF;
                    escaped_medium = !mei_valid(mei); active_medium = mei_valid(mei);
                    if (active_medium) {
                        total_dist += mei.t; ray.o = mei.p; si_t = si_t - mei.t;
                        if (is_spectral) transmittance *= mei.sigma_n; else transmittance *= mei.sigma_n / mei.combined_extinction;
                    }
                }
                if (active_surface && needs_intersection) { cnt.rays++; h = bvh_trace<false, false>(sc, ray); si_t = h.t; needs_intersection = false; }
                active_surface |= escaped_medium;
                if (active_surface) total_dist += si_t;
                active_surface &= (si_t != LR_INF) && !active_medium;
                SurfaceInteraction si;
                if (active_surface) {
                    h.t = si_t; si = compute_si(sc, ray, h, sc.tris_gid);
                    BSDF bsdf(sc, si);
                    transmittance *= bsdf.eval_null_transmission();
                    Ray r2 = spawn_ray(si.p, si.n, ray.d); ray.o = r2.o;
                }
                needs_intersection |= active_surface;
                active &= (active_medium || active_surface) && any_nonzero(transmittance);
                if (active_surface && mesh_is_transition(sc, si.mesh)) medium = target_medium(sc, si, ray.d);
            }
            if (!active) state = 0;
        }
        if (exhausted && !__any_sync(0xffffffffu, state != -2)) break;
    }
    flush_counters(cnt, 0, a.stats);
}

// counters of the queues the second half of an iteration fills (after SHADE_B has consumed the scatter / surface queues)
// `items`: per-stage totals of queue entries consumed (lr_stage_profile.items; index = the host's stage enum, passed in `st`)
struct WVStageIds { int track, shade_a_med, shade_a_surf, escape, trace_s, nee, shade_b_med, shade_b_surf, trace_c, regen; };
__global__ void k_wv_mid_iteration(WVQueues *q, unsigned long long *items, const WVStageIds st) {
    uint32_t surf = 0; for (int b = 0; b < WF_N_SURF; ++b) surf += q->n_surf[b];
    atomicAdd(items + st.track, (unsigned long long) q->n_track); atomicAdd(items + st.shade_a_med, (unsigned long long) q->n_scatter);
    atomicAdd(items + st.shade_b_med, (unsigned long long) q->n_scatter); atomicAdd(items + st.shade_a_surf, (unsigned long long) surf);
    atomicAdd(items + st.shade_b_surf, (unsigned long long) surf); atomicAdd(items + st.escape, (unsigned long long) q->n_escape);
    atomicAdd(items + st.trace_s, (unsigned long long) q->n_shadow); atomicAdd(items + st.nee, (unsigned long long) q->n_nee);
    q->n_track = 0; q->n_scatter = 0; q->n_escape = 0; q->n_shadow = 0; q->n_nee = 0;
    for (int b = 0; b < WF_N_SURF; ++b) q->n_surf[b] = 0;
    q->cursor[0] = q->cursor[1] = q->cursor[2] = 0; q->cursor_nee = 0;
}
// end of an iteration: what is still alive; the consumed trace / regen queues are emptied
__global__ void k_wv_end_iteration(WVQueues *q, uint32_t *live_out, const unsigned long long *lane_next, unsigned long long *lane_next_out,
                                   unsigned long long *items, const WVStageIds st) {
    atomicAdd(items + st.trace_c, (unsigned long long) q->n_trace); atomicAdd(items + st.regen, (unsigned long long) q->n_regen);
    uint32_t live = q->n_track + q->n_escape;
    for (int b = 0; b < WF_N_SURF; ++b) live += q->n_surf[b];
    q->live = live; q->n_trace = 0; q->n_regen = 0;
    *live_out = live; *lane_next_out = *lane_next;
}
__global__ void k_wv_init(WVQueues *q) {
    q->n_trace = 0; q->n_track = q->n_scatter = q->n_escape = q->n_shadow = q->n_nee = q->n_regen = 0;
    for (int b = 0; b < WF_N_SURF; ++b) q->n_surf[b] = 0;
    q->cursor[0] = q->cursor[1] = q->cursor[2] = 0; q->cursor_nee = 0; q->n_alloc = 0; q->live = 0;
}

} // namespace lr
