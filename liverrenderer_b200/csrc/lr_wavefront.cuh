// lr_wavefront.cuh -- wavefront `volpath`: the same per-lane algorithm as sample_volpath()
// (lr_integrators.cuh; reference src/integrators/volpath.cpp:93-554), cut at the points where lanes diverge
// into stage kernels that exchange path state through HBM:
//
//   GEN      seed / continue a lane, camera ray, primary closest hit                 (regen queue)
//   TRACK    russian roulette + delta-tracking loop (null collisions stay in registers) until an EVENT:
//            real scatter -> scatter queue | surface / escape -> surface queue | end of path -> regen queue
//   SHADE_A  scatter: NEE set-up through the phase function; surface: emitter hit + MIS, NEE set-up through the BSDF
//   NEE      ratio-tracking transmittance walk of the shadow segment (volpath.cpp:454-551)
//   SHADE_B  phase / BSDF sampling, next ray + closest hit, medium transition -> next active queue
//
// Every lane keeps its own PCG32 stream (keyed by the GLOBAL lane index) and draws in exactly the order of the
// reference, so a wavefront render is lane-for-lane identical to the per-lane kernel and to the oracle.
// Queues are index arrays compacted with warp ballots (one atomicAdd per warp); surface events are bucketed
// by BSDF class so a warp shades one material model.
#pragma once
#include "lr_integrators.cuh"

// resident 256-thread blocks per SM the stage kernels are compiled for (register cap = 65536 / (256 * LR_WF_OCC))
#ifndef LR_WF_OCC
#define LR_WF_OCC 4
#endif
// resident 128-thread blocks per SM of the two surface-shading kernels (register cap = 65536 / (128 * LR_WF_OCC_SURF))
#ifndef LR_WF_OCC_SURF
#define LR_WF_OCC_SURF (2 * LR_WF_OCC)
#endif

namespace lr {

enum : uint32_t {
    WF_CHANNEL_MASK = 3u, WF_VALID_RAY = 1u << 2, WF_SPEC_CHAIN = 1u << 3, WF_NEE_ACTIVE = 1u << 4, WF_DEAD = 1u << 5,
    WF_NEE_NEEDS_ISECT = 1u << 6, WF_HAS_LANE = 1u << 7,
    WF_MEDIUM_SHIFT = 8, WF_MEDIUM_MASK = 0xffu << 8,          // 0xff: no medium
    WF_PASS_SHIFT = 16                                          // pass index of a multi-pass render (integrator.cpp:279-293)
};
#define WF_N_SURF 3                                             // surface buckets: 0 diffuse/null/dielectric, 1 roughdielectric, 2 principled

// One path slot, 256 B, 16-byte fields: a thread moves its slot with 128-bit loads/stores (full 32 B sectors).
struct __align__(16) WState {
    unsigned long long rng_state, rng_inc;
    float4 o_maxt;          // ray.o, ray.maxt
    float4 d_eta;           // ray.d, eta
    float4 thr_lpdf;        // throughput, last_scatter_direction_pdf
    float4 res_depth;       // result, depth (bits)
    float4 hit;             // t, prim (bits), u, v of the closest hit along the current ray
    float4 lsp_flags;       // last_scatter_p, flags (bits)
    float4 lane_pos;        // lane lo/hi (bits), film sample position
    // ---- NEE walk (volpath.cpp:400-554) ----
    float4 no_maxd;         // shadow ray origin, max_dist
    float4 nd_total;        // ds.d, total_dist
    float4 ntr_mis;         // transmittance, MIS weight
    float4 nw_med;          // throughput * (phase | bsdf) value, walk medium (bits)
    float4 ne_dist;         // emitter_val, ds.dist
    float4 nhit;            // closest hit along the shadow ray
    float4 pad0, pad1;
};
static_assert(sizeof(WState) == 256, "WState layout");

struct WQueues {                       // device-side counters, one 64-byte block
    uint32_t n_active[2]; uint32_t n_scatter; uint32_t n_surf[WF_N_SURF]; uint32_t n_regen; uint32_t live;
    unsigned long long lane_next; uint32_t n_nee; uint32_t cursor; uint32_t pad[4];   // cursor: work cursor of k_wf_track_dda
};

struct WFArgs {
    WState *st; WQueues *q; uint32_t *q_active[2]; uint32_t *q_scatter; uint32_t *q_surf[WF_N_SURF]; uint32_t *q_regen; uint32_t *q_nee;   // q_nee: slots whose NEE walk is live
    uint32_t cap; int cur;            // slots; index of the input active queue
    const DTri *tris_gid; float *film; float *lanes_out;
    uint64_t lane_begin, lane_end; uint32_t seed, spp_per_pass, n_passes;
    unsigned long long *stats;
};

// ---- queue push: ballot + popc, one atomic per warp; call at a point where the whole warp is converged ----
LR_DEV void wf_push(uint32_t *queue, uint32_t *count, uint32_t slot, bool pred) {
    unsigned m = __ballot_sync(0xffffffffu, pred);
    if (m == 0) return;
    unsigned lane = threadIdx.x & 31; uint32_t base = 0;
    if (lane == (unsigned) (__ffs(m) - 1)) base = atomicAdd(count, (uint32_t) __popc(m));
    base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
    if (pred) queue[base + __popc(m & ((1u << lane) - 1u))] = slot;
}

LR_DEV uint32_t wf_flags(const WState &S) { return __float_as_uint(S.lsp_flags.w); }
LR_DEV int wf_medium(uint32_t f) { uint32_t m = (f & WF_MEDIUM_MASK) >> WF_MEDIUM_SHIFT; return m == 0xffu ? -1 : (int) m; }
LR_DEV uint32_t wf_set_medium(uint32_t f, int m) { return (f & ~WF_MEDIUM_MASK) | (((uint32_t) (m < 0 ? 0xff : m)) << WF_MEDIUM_SHIFT); }
LR_DEV Hit wf_hit(float4 h) { Hit r; r.t = h.x; r.prim = __float_as_uint(h.y); r.u = h.z; r.v = h.w; return r; }
LR_DEV float4 wf_hit4(const Hit &h) { return make_float4(h.t, __uint_as_float(h.prim), h.u, h.v); }
LR_DEV int wf_surf_bucket(const DScene &sc, const DTri *tris_gid, const Hit &h) {
    if (!hit_valid(h)) return 0;
    uint32_t k = sc.bsdfs[sc.meshes[__float_as_uint(tris_gid[h.prim].v1.w)].bsdf].kind;
    return k == LR_BSDF_ROUGHDIELECTRIC ? 1 : (k == LR_BSDF_PRINCIPLED ? 2 : 0);
}

// end of a path: splat (ImageBlock::put) or per-lane output
template <int C, bool TO_FILM>
LR_DEV void wf_finish(const DScene &sc, const WFArgs &a, const WState &S, V3 L, uint32_t flags) {
    bool valid = (flags & WF_VALID_RAY) != 0;
    if (TO_FILM) {
        float aovs[C]; aovs[0] = L.x; aovs[1] = L.y; aovs[2] = L.z;
        if (C == 5) { aovs[3] = valid ? 1.f : 0.f; aovs[4] = 1.f; } else aovs[3] = 1.f;
        film_put<C>(sc, a.film, mk2(S.lane_pos.z, S.lane_pos.w), aovs, a.spp_per_pass >= 4);
    } else {
        uint64_t lane = ((uint64_t) __float_as_uint(S.lane_pos.y) << 32) | __float_as_uint(S.lane_pos.x);
        uint32_t pass = flags >> WF_PASS_SHIFT;
        reinterpret_cast<float4 *>(a.lanes_out)[(lane - a.lane_begin) * a.n_passes + pass] = make_float4(L.x, L.y, L.z, valid ? 1.f : 0.f);
    }
}

#define WF_FOR_QUEUE(COUNT)                                                                             \
    const uint32_t _n = (COUNT);                                                                        \
    const uint32_t _stride = gridDim.x * blockDim.x;                                                    \
    for (uint32_t _base = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u; _base < _n; _base += _stride) \
        if (const uint32_t qi = _base + (threadIdx.x & 31u); true)

// ------------------------------------------------------------------------------------------
// GEN: every free slot takes the next pass of its lane or the next global lane (integrator.cpp:321-356)
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wf_gen(const __grid_constant__ DScene sc, const __grid_constant__ WFArgs a) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0; unsigned long long nsamples = 0;
    uint32_t *q_out = a.q_active[a.cur ^ 1], *n_out = &a.q->n_active[a.cur ^ 1];
    WF_FOR_QUEUE(a.q->n_regen) {
        bool item = qi < _n; uint32_t slot = item ? a.q_regen[qi] : 0; bool live = false;
        WState *S = a.st + slot;
        Sampler smp; uint64_t lane = 0; uint32_t pass = 0; bool need_lane = item;
        if (item) {
            uint32_t f = wf_flags(*S);
            if ((f & WF_HAS_LANE) && (f >> WF_PASS_SHIFT) + 1 < a.n_passes) {          // same lane, next pass: the stream continues
                need_lane = false; live = true; pass = (f >> WF_PASS_SHIFT) + 1;
                smp.state = S->rng_state; smp.inc = S->rng_inc; smp.advance(a.spp_per_pass);
                lane = ((uint64_t) __float_as_uint(S->lane_pos.y) << 32) | __float_as_uint(S->lane_pos.x);
            }
        }
        {   // hand out fresh global lanes, one atomic per warp
            unsigned m = __ballot_sync(0xffffffffu, need_lane); unsigned l = threadIdx.x & 31; unsigned long long base = 0;
            if (m) {
                if (l == (unsigned) (__ffs(m) - 1)) base = atomicAdd(&a.q->lane_next, (unsigned long long) __popc(m));
                base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
                if (need_lane) {
                    lane = a.lane_begin + base + __popc(m & ((1u << l) - 1u));
                    if (lane < a.lane_end) { live = true; seed_sampler(sc, smp, a.seed, lane, a.spp_per_pass, a.n_passes); }
                }
            }
        }
        if (item && !live) { S->lsp_flags.w = __uint_as_float(WF_DEAD); }
        if (live) {
            LaneSetup L; L.smp = smp; lane_begin_sample(sc, L, lane, a.spp_per_pass); lane_next_ray(sc, L);
            uint32_t channel = (uint32_t) fminf(L.smp.next_1d() * 3.f, 2.f);                // volpath.cpp:117-120
            uint32_t f = channel | WF_HAS_LANE | (pass << WF_PASS_SHIFT);
            if (!sc.hide_emitters && sc.env >= 0) f |= WF_VALID_RAY;                         // :103
            if (!sc.hide_emitters) f |= WF_SPEC_CHAIN;                                       // :114
            f = wf_set_medium(f, sc.sensor.medium);
            cnt.rays++; Hit h = bvh_trace<false, false>(sc, L.ray); nsamples++;
            bool box = sc.sensor.rfilter == LR_RFILTER_BOX; V2 fp = box ? L.pos : L.sample_pos;
            S->rng_state = L.smp.state; S->rng_inc = L.smp.inc;
            S->o_maxt = make_float4(L.ray.o.x, L.ray.o.y, L.ray.o.z, L.ray.maxt); S->d_eta = make_float4(L.ray.d.x, L.ray.d.y, L.ray.d.z, 1.f);
            S->thr_lpdf = make_float4(1.f, 1.f, 1.f, 1.f); S->res_depth = make_float4(0.f, 0.f, 0.f, __uint_as_float(0u));
            S->hit = wf_hit4(h); S->lsp_flags = make_float4(0.f, 0.f, 0.f, __uint_as_float(f));
            S->lane_pos = make_float4(__uint_as_float((uint32_t) lane), __uint_as_float((uint32_t) (lane >> 32)), fp.x, fp.y);
        }
        wf_push(q_out, n_out, slot, live);
    }
    flush_counters(cnt, nsamples, a.stats);
}

// ------------------------------------------------------------------------------------------
// TRACK: volpath.cpp:160-253 -- russian roulette, free-flight sampling, null collisions in registers
// ------------------------------------------------------------------------------------------
template <int C, bool TO_FILM>
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wf_track(const __grid_constant__ DScene sc, const __grid_constant__ WFArgs a) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    const uint32_t *q_in = a.q_active[a.cur];
    WF_FOR_QUEUE(a.q->n_active[a.cur]) {
        bool item = qi < _n; uint32_t slot = item ? q_in[qi] : 0; WState *S = a.st + slot;
        int event = -1;                                   // -1 none, 0 end of path, 1 scatter, 2.. surface bucket
        if (item) {
            Sampler smp; smp.state = S->rng_state; smp.inc = S->rng_inc;
            float4 t4 = S->o_maxt; Ray ray; ray.o = mk3(t4.x, t4.y, t4.z); ray.maxt = t4.w;
            t4 = S->d_eta; ray.d = mk3(t4.x, t4.y, t4.z); float eta = t4.w;
            t4 = S->thr_lpdf; V3 thr = mk3(t4.x, t4.y, t4.z); float lpdf = t4.w;
            float4 h4 = S->hit; float hit_t = h4.x;
            float4 lf = S->lsp_flags; uint32_t f = __float_as_uint(lf.w); V3 lsp = mk3(lf.x, lf.y, lf.z);
            uint32_t depth = __float_as_uint(S->res_depth.w);
            const uint32_t channel = f & WF_CHANNEL_MASK; const int medium = wf_medium(f);
            for (;;) {
                bool active = any_nonzero(thr);
                float q = fminf(hmax(thr) * sqr(eta), .95f);
                bool perform_rr = depth > (uint32_t) sc.rr_depth;
                if (active) { float u = smp.next_1d(); active &= (u < q) || !perform_rr; }
                if (perform_rr) thr *= 1.f / q;
                active &= depth < sc.max_depth;
                if (!active) { event = 0; break; }
                if (medium >= 0 && medium_uses_dda(sc, sc.media[medium])) {
                    // production tracking: delta tracking against the local majorants of the super-grid.  The majorant is one
                    // scalar for all channels (heterogeneous.cpp:163), so Tr/pdf reduces to 1/mu at a collision and 1 at a surface.
                    const DMedium &m = sc.media[medium]; const DGrid &g = sc.grids[m.m.sigma_t_grid];
                    float mint, maxt;
                    if (medium_clip(g, ray, mint, maxt)) {
                        const float t_end = fminf(maxt, hit_t);
                        if (mint < t_end) {
                            DDA dd; dda_init(g, ray, mint, dd); float t = mint, tau = -logf(1.f - smp.next_1d()), mu;
                            while (dda_walk(g, m.m.scale, dd, t, tau, t_end, mu)) {
                                cnt.collisions++;
                                V3 p = ray_at(ray, t);
                                V3 st = grid_eval<false>(g, p, nullptr) * m.m.scale, sn = mk3(mu) - st;
                                float pn = hmean(sn / mu);
                                if (smp.next_1d() < pn) { thr *= sn / (mu * pn); tau = -logf(1.f - smp.next_1d()); continue; }
                                depth += 1; lsp = p; cnt.real_scatters++;
                                if (!(depth < sc.max_depth)) { event = 0; break; }
                                thr *= (st * medium_albedo(sc, m, p)) / (mu * hmean(st / mu));
                                ray.o = p; hit_t = hit_t - t; event = 1; break;
                            }
                        }
                    }
                    if (event < 0) event = 2;
                    break;
                }
                if (medium >= 0) {
                    const DMedium &m = sc.media[medium];
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
                        if (smp.next_1d() < null_scatter_prob) {                       // null collision: stay in the loop
                            if (is_spectral) thr *= mei.sigma_n / null_scatter_prob;
                            ray.o = mei.p; hit_t = hit_t - mei.t;
                            continue;
                        }
                        depth += 1; lsp = mei.p; cnt.real_scatters++;
                        if (!(depth < sc.max_depth)) { event = 0; break; }
                        if (is_spectral) thr *= mei.sigma_s / hmean(mei.sigma_t / mei.combined_extinction);
                        else thr *= mei.sigma_s / mei.sigma_t;
                        ray.o = mei.p; event = 1; break;
                    }
                }
                event = 2; break;                          // surface interaction or escape (bucket chosen below)
            }
            if (event == 0) {
                float4 r4 = S->res_depth; wf_finish<C, TO_FILM>(sc, a, *S, mk3(r4.x, r4.y, r4.z), f);
                S->rng_state = smp.state; S->rng_inc = smp.inc;        // a later pass of this lane continues the stream
            } else {
                if (event == 2) event = 2 + wf_surf_bucket(sc, a.tris_gid, wf_hit(h4));
                S->rng_state = smp.state; S->rng_inc = smp.inc;
                S->o_maxt = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.maxt);
                S->thr_lpdf = make_float4(thr.x, thr.y, thr.z, lpdf);
                S->hit.x = hit_t; S->res_depth.w = __uint_as_float(depth);
                S->lsp_flags = make_float4(lsp.x, lsp.y, lsp.z, __uint_as_float(f));
            }
        }
        wf_push(a.q_regen, &a.q->n_regen, slot, event == 0);
        wf_push(a.q_scatter, &a.q->n_scatter, slot, event == 1);
#pragma unroll
        for (int b = 0; b < WF_N_SURF; ++b) wf_push(a.q_surf[b], &a.q->n_surf[b], slot, event == 2 + b);
    }
    flush_counters(cnt, 0, a.stats);
}

// TRACK for scenes whose media all use the majorant-DDA (production mode): the same per-item work as k_wf_track, run as a
// persistent "while-while" loop.  The number of super-voxels a free flight crosses is roughly exponentially distributed, so a
// warp that waits for its longest walk keeps 4-5 of 32 lanes busy (ncu source view of k_wf_track: lr_medium.cuh:101-118 at 4.5
// threads per instruction).  Here a lane that has finished its item writes it out and, once half of the warp is idle, the idle
// lanes pull the next queue entries (one atomicAdd on q->cursor per refill), so the walk loop stays populated.  Per-lane RNG
// draw order is unchanged: RR, first optical depth, then (null test, next optical depth) per tentative collision.
#ifndef LR_TRACK_REFILL
#define LR_TRACK_REFILL 16            // refill when at most this many lanes of the warp are still walking
#endif
#ifndef LR_TRACK_STEPS
#define LR_TRACK_STEPS 6              // super-voxel steps per lane between two converged points
#endif
template <int C, bool TO_FILM>
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wf_track_dda(const __grid_constant__ DScene sc, const __grid_constant__ WFArgs a) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    const uint32_t *q_in = a.q_active[a.cur]; const uint32_t n_items = a.q->n_active[a.cur];
    const unsigned lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    uint32_t slot = 0, depth = 0, f = 0; int event = -2;                  // -2 idle, -1 walking, 0 end of path, 1 scatter, 2 surface
    Sampler smp; smp.state = smp.inc = 0; Ray ray; ray.o = ray.d = mk3(0.f); ray.maxt = 0.f; V3 thr = mk3(0.f);
    float hit_t = 0.f, t = 0.f, tau = 0.f, t_end = 0.f, mu_scale = 0.f; int grid_id = 0; DDA dd{};
    bool exhausted = false;
    for (;;) {
        const unsigned walking = __ballot_sync(0xffffffffu, event == -1);
        if (!exhausted && __popc(walking) <= LR_TRACK_REFILL) {
            const unsigned idle = ~walking; uint32_t base = 0;
            if (lane == (unsigned) (__ffs(idle) - 1)) base = atomicAdd(&a.q->cursor, (uint32_t) __popc(idle));
            base = __shfl_sync(0xffffffffu, base, __ffs(idle) - 1);
            exhausted = base + (uint32_t) __popc(idle) >= n_items;
            const uint32_t qi = base + (uint32_t) __popc(idle & lt);
            if (event == -2 && qi < n_items) {
                slot = q_in[qi]; const WState *S = a.st + slot;
                smp.state = S->rng_state; smp.inc = S->rng_inc;
                float4 t4 = S->o_maxt; ray.o = mk3(t4.x, t4.y, t4.z); ray.maxt = t4.w;
                t4 = S->d_eta; ray.d = mk3(t4.x, t4.y, t4.z); const float eta = t4.w;
                t4 = S->thr_lpdf; thr = mk3(t4.x, t4.y, t4.z);
                hit_t = S->hit.x; f = __float_as_uint(S->lsp_flags.w); depth = __float_as_uint(S->res_depth.w);
                const int medium = wf_medium(f);
                // volpath.cpp:160-178: russian roulette, depth test
                bool active = any_nonzero(thr);
                float q = fminf(hmax(thr) * sqr(eta), .95f);
                bool perform_rr = depth > (uint32_t) sc.rr_depth;
                if (active) { float u = smp.next_1d(); active &= (u < q) || !perform_rr; }
                if (perform_rr) thr *= 1.f / q;
                active &= depth < sc.max_depth;
                event = active ? 2 : 0;
                if (active && medium >= 0) {
                    const DMedium &m = sc.media[medium]; grid_id = m.m.sigma_t_grid; mu_scale = m.m.scale; const DGrid &g = sc.grids[grid_id];
                    float mint, maxt;
                    if (medium_clip(g, ray, mint, maxt)) {
                        t_end = fminf(maxt, hit_t);
                        if (mint < t_end) { dda_init(g, ray, mint, dd); t = mint; tau = -logf(1.f - smp.next_1d()); event = -1; }
                    }
                }
            }
        }
        // ---- walk: bring every walking lane to its next tentative collision (or the end of its segment) ----
        bool collide = false; float mu = 0.f;
        if (event == -1) {
            const DGrid &g = sc.grids[grid_id];
#pragma unroll 1
            for (int k = 0; k < LR_TRACK_STEPS; ++k) {
                mu = mu_scale * dda_majorant(g, dd);
                const float t_exit = fminf(fminf(dd.tnx, dd.tny), fminf(dd.tnz, t_end));
                const float dtau = mu * (t_exit - t);
                if (dtau >= tau && mu > 0.f) { t = fminf(t + tau / mu, t_exit); collide = true; break; }
                tau -= dtau; t = t_exit;
                if (t_exit >= t_end) { event = 2; break; }
                dda_advance(g, dd);
            }
        }
        // ---- tentative collisions, evaluated together ----
        if (collide) {
            const DGrid &g = sc.grids[grid_id]; const DMedium &m = sc.media[wf_medium(f)];
            cnt.collisions++;
            V3 p = ray_at(ray, t);
            V3 st = grid_eval<false>(g, p, nullptr) * mu_scale, sn = mk3(mu) - st;
            float pn = hmean(sn / mu);
            if (smp.next_1d() < pn) { thr *= sn / (mu * pn); tau = -logf(1.f - smp.next_1d()); }
            else {
                depth += 1; cnt.real_scatters++;
                WState *S = a.st + slot; S->lsp_flags.x = p.x; S->lsp_flags.y = p.y; S->lsp_flags.z = p.z;
                if (!(depth < sc.max_depth)) event = 0;
                else { thr *= (st * medium_albedo(sc, m, p)) / (mu * hmean(st / mu)); ray.o = p; hit_t = hit_t - t; event = 1; }
            }
        }
        // ---- write finished items out and queue them (converged: wf_push uses warp ballots) ----
        const bool fin = event >= 0;
        if (__any_sync(0xffffffffu, fin)) {
            WState *S = a.st + slot;
            if (fin) {
                S->rng_state = smp.state; S->rng_inc = smp.inc;            // a later pass of this lane continues the stream
                if (event == 0) { float4 r4 = S->res_depth; wf_finish<C, TO_FILM>(sc, a, *S, mk3(r4.x, r4.y, r4.z), f); }
                else {
                    if (event == 2) event = 2 + wf_surf_bucket(sc, a.tris_gid, wf_hit(S->hit));
                    S->o_maxt = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.maxt);
                    S->thr_lpdf.x = thr.x; S->thr_lpdf.y = thr.y; S->thr_lpdf.z = thr.z;
                    S->hit.x = hit_t; S->res_depth.w = __uint_as_float(depth);
                }
            }
            wf_push(a.q_regen, &a.q->n_regen, slot, event == 0);
            wf_push(a.q_scatter, &a.q->n_scatter, slot, event == 1);
#pragma unroll
            for (int b = 0; b < WF_N_SURF; ++b) wf_push(a.q_surf[b], &a.q->n_surf[b], slot, event == 2 + b);
            if (fin) event = -2;
        }
        if (exhausted && !__any_sync(0xffffffffu, event == -1)) break;
    }
    flush_counters(cnt, 0, a.stats);
}

// NEE set-up shared by both SHADE_A kernels: Scene::sample_emitter_direction + the first lines of
// VolumetricPathIntegrator::sample_emitter (volpath.cpp:400-452), plus the shadow ray's closest hit.
LR_DEV bool wf_nee_setup(const DScene &sc, const WFArgs &a, WState *S, Sampler &smp, V3 ref_p, V3 ref_n, const SurfaceInteraction *ref_si,
                         int medium, DirectionSample &ds, Counters &cnt) {
    V3 emitter_val; ds = sample_emitter_direction(sc, ref_p, smp.next_2d(), emitter_val);
    if (ds.pdf == 0.f) return false;
    Ray ray = spawn_ray_to(ref_p, ref_n, ds.p);
    if (ref_si && mesh_is_transition(sc, ref_si->mesh)) medium = target_medium(sc, *ref_si, ray.d);
    cnt.rays++; Hit h;
    // production mode without index-matched boundaries: every surface blocks a shadow ray, so the first accepted triangle decides
    // (any-hit query, no closest hit needed); a miss leaves the whole segment to the medium walk
    if (sc.integ.tracking == LR_TRACKING_DDA && !sc.any_null_bsdf) { h = bvh_trace<true, false>(sc, ray); if (hit_valid(h)) return false; }
    else h = bvh_trace<false, false>(sc, ray);
    if (sc.integ.tracking == LR_TRACKING_DDA && hit_valid(h)) {
        // production mode: opaque-occluder pre-test.  A non-null surface anywhere on the shadow segment makes the transmittance
        // exactly 0 (eval_null_transmission), so the ratio-tracking walk is skipped; only the RNG consumption differs.
        bool blocked = true;
        if (sc.any_null_bsdf) { cnt.rays++; blocked = hit_valid(bvh_trace<true, true>(sc, ray)); }
        if (blocked) return false;
    }
    S->no_maxd = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.maxt);
    S->nd_total = make_float4(ray.d.x, ray.d.y, ray.d.z, 0.f);
    S->ne_dist = make_float4(emitter_val.x, emitter_val.y, emitter_val.z, ds.dist);
    S->nhit = wf_hit4(h);
    S->nw_med.w = __int_as_float(medium);
    return true;
}

// SHADE_A, medium scatter (volpath.cpp:256-283)
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wf_shade_a_medium(const __grid_constant__ DScene sc, const __grid_constant__ WFArgs a) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    WF_FOR_QUEUE(a.q->n_scatter) {
        bool item = qi < _n, nee = false; uint32_t slot = item ? a.q_scatter[qi] : 0;
        if (item) {
            WState *S = a.st + slot;
            Sampler smp; smp.state = S->rng_state; smp.inc = S->rng_inc;
            uint32_t f = wf_flags(*S); int medium = wf_medium(f);
            const DMedium &m = sc.media[medium]; Phase phase; phase.kind = m.m.phase; phase.g = m.m.phase_g;
            bool sample_emitters = m.m.sample_emitters != 0;
            f |= WF_VALID_RAY; f = sample_emitters ? (f & ~WF_SPEC_CHAIN) : (f | WF_SPEC_CHAIN); f &= ~(WF_NEE_ACTIVE | WF_DEAD);
            if (sample_emitters) {
                float4 o = S->o_maxt, d = S->d_eta, t = S->thr_lpdf; V3 p = mk3(o.x, o.y, o.z), wi = -mk3(d.x, d.y, d.z), thr = mk3(t.x, t.y, t.z);
                DirectionSample ds;
                if (wf_nee_setup(sc, a, S, smp, p, mk3(0.f), nullptr, medium, ds, cnt)) {
                    float phase_val = phase.eval_pdf(wi, ds.d); V3 w = thr * phase_val;
                    S->ntr_mis = make_float4(1.f, 1.f, 1.f, mis_weight(ds.pdf, ds.delta ? 0.f : phase_val));
                    S->nw_med.x = w.x; S->nw_med.y = w.y; S->nw_med.z = w.z;
                    f |= WF_NEE_ACTIVE;
                }
                S->rng_state = smp.state; S->rng_inc = smp.inc;
            }
            S->lsp_flags.w = __uint_as_float(f); nee = (f & WF_NEE_ACTIVE) != 0;
        }
        wf_push(a.q_nee, &a.q->n_nee, slot, nee);
    }
    flush_counters(cnt, 0, a.stats);
}

// SHADE_A, surface / escape (volpath.cpp:300-362)
template <int C, bool TO_FILM>
__global__ void __launch_bounds__(128, LR_WF_OCC_SURF) k_wf_shade_a_surface(const __grid_constant__ DScene sc, const __grid_constant__ WFArgs a, int bucket) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    WF_FOR_QUEUE(a.q->n_surf[bucket]) {
        bool item = qi < _n; uint32_t slot = item ? a.q_surf[bucket][qi] : 0; bool dead = false, nee = false;
        if (item) {
            WState *S = a.st + slot;
            Sampler smp; smp.state = S->rng_state; smp.inc = S->rng_inc;
            uint32_t f = wf_flags(*S) & ~(WF_NEE_ACTIVE | WF_DEAD); int medium = wf_medium(f);
            float4 o = S->o_maxt, d = S->d_eta, t = S->thr_lpdf, r4 = S->res_depth;
            Ray ray; ray.o = mk3(o.x, o.y, o.z); ray.d = mk3(d.x, d.y, d.z); ray.maxt = o.w;
            V3 thr = mk3(t.x, t.y, t.z), result = mk3(r4.x, r4.y, r4.z); uint32_t depth = __float_as_uint(r4.w);
            Hit h = wf_hit(S->hit);
            SurfaceInteraction si = compute_si(sc, ray, h, a.tris_gid);
            // the segment started outside any medium <=> the reference intersected in this very iteration (`intersect`, :301-303)
            if (sc.hide_emitters && medium < 0 && depth == 0u && si_valid(si) && sc.meshes[si.mesh].emitter >= 0) {        // :307-323
                Ray r2 = spawn_ray(si.p, si.n, ray.d); Hit pi = skip_area_emitters(sc, a.tris_gid, r2, cnt);
                si = compute_si(sc, r2, pi, a.tris_gid); h = pi; S->hit = wf_hit4(h);
                S->o_maxt = make_float4(r2.o.x, r2.o.y, r2.o.z, r2.maxt);       // SHADE_B recomputes si from (ray, hit): keep them consistent
            }
            bool count_direct = depth == 0u || (f & WF_SPEC_CHAIN);
            int emitter = emitter_of(sc, si);
            if (emitter >= 0 && !(depth == 0u && sc.hide_emitters)) {                                                     // :326-345
                float emitter_pdf = 1.f;
                if (!count_direct) { float4 l = S->lsp_flags; DirectionSample ds = ds_from_si(sc, si, mk3(l.x, l.y, l.z)); emitter_pdf = pdf_emitter_direction(sc, ds); }
                V3 emitted = eval_emitter_at(sc, emitter, si);
                result += count_direct ? thr * emitted : thr * mis_weight(t.w, emitter_pdf) * emitted;
            }
            if (!si_valid(si)) {                                   // escaped to the environment: the path ends here
                dead = true; f |= WF_DEAD; wf_finish<C, TO_FILM>(sc, a, *S, result, f);
            } else {
                BSDF bsdf(sc, si);
                if ((bsdf.flags() & F_SMOOTH) && (depth + 1 < sc.max_depth)) {                                           // :350-362
                    DirectionSample ds;
                    if (wf_nee_setup(sc, a, S, smp, si.p, si.n, &si, medium, ds, cnt)) {
                        V3 wo = si.sh.to_local(ds.d); float bsdf_pdf; V3 bsdf_val = bsdf.eval_pdf(si.wi, wo, bsdf_pdf);
                        V3 w = thr * bsdf_val;
                        S->ntr_mis = make_float4(1.f, 1.f, 1.f, mis_weight(ds.pdf, ds.delta ? 0.f : bsdf_pdf));
                        S->nw_med.x = w.x; S->nw_med.y = w.y; S->nw_med.z = w.z;
                        f |= WF_NEE_ACTIVE;
                    }
                }
            }
            S->rng_state = smp.state; S->rng_inc = smp.inc;
            S->res_depth = make_float4(result.x, result.y, result.z, r4.w);
            S->lsp_flags.w = __uint_as_float(f); nee = (f & WF_NEE_ACTIVE) != 0;
        }
        wf_push(a.q_regen, &a.q->n_regen, slot, dead);
        wf_push(a.q_nee, &a.q->n_nee, slot, nee);
    }
    flush_counters(cnt, 0, a.stats);
}

// ------------------------------------------------------------------------------------------
// NEE: transmittance along the shadow segment (volpath.cpp:454-551), ratio tracking through media and
// index-matched (null) boundaries.  Runs over the NEE queue the SHADE_A kernels fill.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wf_nee(const __grid_constant__ DScene sc, const __grid_constant__ WFArgs a, uint32_t surf_mask) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0; (void) surf_mask;
    {
    const uint32_t *queue = a.q_nee;
    WF_FOR_QUEUE(a.q->n_nee) {
        if (qi < _n) {
            WState *S = a.st + queue[qi]; uint32_t f = wf_flags(*S);
            if ((f & WF_NEE_ACTIVE) && !(f & WF_DEAD)) {
                Sampler smp; smp.state = S->rng_state; smp.inc = S->rng_inc;
                const uint32_t channel = f & WF_CHANNEL_MASK;
                float4 o = S->no_maxd, d = S->nd_total; Ray ray; ray.o = mk3(o.x, o.y, o.z); ray.d = mk3(d.x, d.y, d.z);
                const float max_dist = o.w, ds_dist = S->ne_dist.w; float total_dist = 0.f;
                int medium = __float_as_int(S->nw_med.w);
                Hit h = wf_hit(S->nhit); float si_t = h.t; bool needs_intersection = false, active = true;
                V3 transmittance = mk3(1.f);
                while (active) {
                    float remaining_dist = max_dist - total_dist; ray.maxt = remaining_dist;
                    active &= remaining_dist > 0.f;
                    if (!active) break;
                    bool escaped_medium = false, active_medium = medium >= 0, active_surface = !active_medium; float mei_t = LR_INF;
                    if (active_medium && medium_uses_dda(sc, sc.media[medium])) {
                        // ratio tracking over the whole medium segment against the local majorants: T *= sigma_n / mu per tentative collision
                        const DMedium &m = sc.media[medium]; const DGrid &g = sc.grids[m.m.sigma_t_grid];
                        if (needs_intersection) { cnt.rays++; h = bvh_trace<false, false>(sc, ray); si_t = h.t; needs_intersection = false; }
                        float mint, maxt;
                        if (medium_clip(g, ray, mint, maxt)) {
                            const float t_end = fminf(maxt, si_t);
                            if (mint < t_end) {
                                DDA dd; dda_init(g, ray, mint, dd); float t = mint, tau = -logf(1.f - smp.next_1d()), mu;
                                while (dda_walk(g, m.m.scale, dd, t, tau, t_end, mu)) {
                                    cnt.collisions++;
                                    V3 st = grid_eval<false>(g, ray_at(ray, t), nullptr) * m.m.scale;
                                    transmittance *= (mk3(mu) - st) / mu;
                                    if (!any_nonzero(transmittance)) break;
                                    tau = -logf(1.f - smp.next_1d());
                                }
                            }
                        }
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
                        mei_t = mei.t;
                    }
                    if (active_surface && needs_intersection) { cnt.rays++; h = bvh_trace<false, false>(sc, ray); si_t = h.t; needs_intersection = false; }
                    active_surface |= escaped_medium;
                    if (active_surface) total_dist += si_t;
                    active_surface &= (si_t != LR_INF) && !active_medium;
                    SurfaceInteraction si;
                    if (active_surface) {
                        h.t = si_t; si = compute_si(sc, ray, h, a.tris_gid);
                        BSDF bsdf(sc, si);
                        transmittance *= bsdf.eval_null_transmission();
                        Ray r2 = spawn_ray(si.p, si.n, ray.d); ray.o = r2.o;
                    }
                    ray.maxt = remaining_dist;
                    needs_intersection |= active_surface;
                    active &= (active_medium || active_surface) && any_nonzero(transmittance);
                    if (active_surface && mesh_is_transition(sc, si.mesh)) medium = target_medium(sc, si, ray.d);
                    (void) mei_t;
                }
                // result += throughput * (phase | bsdf) value * mis * (transmittance * emitter_val)     (:279-283, :356-361)
                float4 w = S->nw_med, e = S->ne_dist, r4 = S->res_depth; float mis = S->ntr_mis.w;
                V3 emitted = transmittance * mk3(e.x, e.y, e.z);
                V3 res = mk3(r4.x, r4.y, r4.z) + mk3(w.x, w.y, w.z) * mis * emitted;
                S->res_depth = make_float4(res.x, res.y, res.z, r4.w);
                S->rng_state = smp.state; S->rng_inc = smp.inc;
            }
        }
    }
    }
    flush_counters(cnt, 0, a.stats);
}

// SHADE_B, medium scatter: phase-function sampling (volpath.cpp:285-297) + closest hit of the new ray
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wf_shade_b_medium(const __grid_constant__ DScene sc, const __grid_constant__ WFArgs a) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    uint32_t *q_out = a.q_active[a.cur ^ 1], *n_out = &a.q->n_active[a.cur ^ 1];
    WF_FOR_QUEUE(a.q->n_scatter) {
        bool item = qi < _n; uint32_t slot = item ? a.q_scatter[qi] : 0;
        if (item) {
            WState *S = a.st + slot;
            Sampler smp; smp.state = S->rng_state; smp.inc = S->rng_inc;
            const DMedium &m = sc.media[wf_medium(wf_flags(*S))]; Phase phase; phase.kind = m.m.phase; phase.g = m.m.phase_g;
            float4 o = S->o_maxt, d = S->d_eta;
            (void) smp.next_1d(); V2 s2 = smp.next_2d();
            float phase_pdf; V3 wo = phase.sample(-mk3(d.x, d.y, d.z), s2, phase_pdf);
            Ray ray; ray.o = mk3(o.x, o.y, o.z); ray.d = mk3(d.x, d.y, d.z); ray.maxt = o.w;
            if (phase_pdf > 0.f) { ray.d = wo; ray.maxt = LR_LARGEST; S->thr_lpdf.w = phase_pdf; }
            cnt.rays++; Hit h = bvh_trace<false, false>(sc, ray);
            S->rng_state = smp.state; S->rng_inc = smp.inc;
            S->o_maxt.w = ray.maxt; S->d_eta = make_float4(ray.d.x, ray.d.y, ray.d.z, d.w); S->hit = wf_hit4(h);
        }
        wf_push(q_out, n_out, slot, item);
    }
    flush_counters(cnt, 0, a.stats);
}

// SHADE_B, surface: BSDF sampling, next ray, medium transition (volpath.cpp:364-389)
__global__ void __launch_bounds__(128, LR_WF_OCC_SURF) k_wf_shade_b_surface(const __grid_constant__ DScene sc, const __grid_constant__ WFArgs a, int bucket) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    uint32_t *q_out = a.q_active[a.cur ^ 1], *n_out = &a.q->n_active[a.cur ^ 1];
    WF_FOR_QUEUE(a.q->n_surf[bucket]) {
        bool item = qi < _n; uint32_t slot = item ? a.q_surf[bucket][qi] : 0; bool live = false;
        if (item) {
            WState *S = a.st + slot; uint32_t f = wf_flags(*S);
            if (!(f & WF_DEAD)) {
                live = true;
                Sampler smp; smp.state = S->rng_state; smp.inc = S->rng_inc;
                float4 o = S->o_maxt, d = S->d_eta, t = S->thr_lpdf;
                Ray ray; ray.o = mk3(o.x, o.y, o.z); ray.d = mk3(d.x, d.y, d.z); ray.maxt = o.w;
                SurfaceInteraction si = compute_si(sc, ray, wf_hit(S->hit), a.tris_gid);
                BSDF bsdf(sc, si);
                float s1 = smp.next_1d(); V2 s2 = smp.next_2d();
                V3 bsdf_val; BSDFSample bs = bsdf.sample(si.wi, s1, s2, bsdf_val);
                V3 thr = mk3(t.x, t.y, t.z) * bsdf_val; float eta = d.w * bs.eta, lpdf = t.w;
                ray = spawn_ray(si.p, si.n, si.sh.to_world(bs.wo));
                bool non_null = !(bs.sampled_type & F_NULL);
                if (non_null) {
                    S->res_depth.w = __uint_as_float(__float_as_uint(S->res_depth.w) + 1u);
                    S->lsp_flags.x = si.p.x; S->lsp_flags.y = si.p.y; S->lsp_flags.z = si.p.z; lpdf = bs.pdf; f |= WF_VALID_RAY;
                }
                bool spec = (f & WF_SPEC_CHAIN) != 0;
                spec |= non_null && (bs.sampled_type & F_DELTA); spec &= !(bs.sampled_type & F_SMOOTH);
                f = spec ? (f | WF_SPEC_CHAIN) : (f & ~WF_SPEC_CHAIN);
                if (mesh_is_transition(sc, si.mesh)) f = wf_set_medium(f, target_medium(sc, si, ray.d));
                cnt.rays++; Hit h = bvh_trace<false, false>(sc, ray);
                S->rng_state = smp.state; S->rng_inc = smp.inc;
                S->o_maxt = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.maxt); S->d_eta = make_float4(ray.d.x, ray.d.y, ray.d.z, eta);
                S->thr_lpdf = make_float4(thr.x, thr.y, thr.z, lpdf); S->hit = wf_hit4(h);
                S->lsp_flags.w = __uint_as_float(f);
            }
        }
        wf_push(q_out, n_out, slot, live);
    }
    flush_counters(cnt, 0, a.stats);
}

__global__ void k_wf_begin_iteration(WQueues *q, int cur) {     // counters of the queues this iteration fills
    q->n_active[cur ^ 1] = 0; q->n_scatter = 0; q->n_regen = 0; q->n_nee = 0; q->cursor = 0;
    for (int b = 0; b < WF_N_SURF; ++b) q->n_surf[b] = 0;
}
__global__ void k_wf_init(WQueues *q, uint32_t *q_regen, uint32_t cap, WState *st) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { q->n_active[0] = q->n_active[1] = 0; q->n_scatter = 0; q->n_regen = cap; q->lane_next = 0; q->n_nee = 0; q->cursor = 0; for (int b = 0; b < WF_N_SURF; ++b) q->n_surf[b] = 0; }
    if (i < cap) { q_regen[i] = i; st[i].lsp_flags.w = __uint_as_float(0u); }
}

} // namespace lr
