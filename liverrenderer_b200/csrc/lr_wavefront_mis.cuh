// lr_wavefront_mis.cuh -- wavefront `volpathmis` (reference: src/integrators/volpathmis.cpp:97-700): MisPath::step (lr_volpathmis.cuh) cut
// at the same places as the volpath wavefront (lr_wavefront.cuh) -- GEN, TRACE_C, TRACK, SHADE_A (medium / surface), ESCAPE, TRACE_S, NEE,
// SHADE_B (medium / surface) -- on its planes, queues, GEN, persistent TRACE stage and multi-batch host loop.  Tracking is the reference's
// own (one global majorant, identical draw order); what differs from volpath is the bookkeeping: the path carries p / f instead of a
// throughput, two 3 x 3 matrices with spectral MIS (one row each without), kept in eight extra 16-byte records per slot (WVArgs::mis):
//   0-2  p_over_f rows          3-5  p_over_f_nee rows
//   6    emitter_val, ds.pdf    7    (phase | bsdf) value, pdf of that direction under the other strategy        (the emitter sample in flight)
// `last_event_was_null` is only ever true between two loop iterations inside TRACK.  Same PCG32 draws in the same order as MisPath::step:
// a wavefront render is lane-for-lane the per-lane kernel's.
#pragma once
#include "lr_wavefront.cuh"
#include "lr_volpathmis.cuh"

namespace lr {

#define LR_MIS_OCC 2                    // 128 registers per thread: two weight matrices (+ two more in NEE) live in registers

template <bool S> LR_DEV MisW<S> misw_load(const float4 *m, int base) {
    MisW<S> w;
#pragma unroll
    for (int i = 0; i < (S ? 3 : 1); ++i) { const float4 v = m[base + i]; w.r[i] = mk3(v.x, v.y, v.z); }
    return w;
}
template <bool S> LR_DEV void misw_store(float4 *m, int base, const MisW<S> &w) {
#pragma unroll
    for (int i = 0; i < (S ? 3 : 1); ++i) m[base + i] = make_float4(w.r[i].x, w.r[i].y, w.r[i].z, 0.f);
}
LR_DEV float4 *mis_rec(const WVArgs &a, uint32_t slot) { return a.mis + (size_t) slot * 8u; }

// volpathmis.cpp:222-234 at the top of a loop iteration (drawn by the stage that spawned the ray; no draw lies in between)
template <bool S>
LR_DEV bool mis_next_iteration(const DScene &sc, Sampler &smp, MisW<S> &P, float eta, uint32_t depth, uint32_t channel, bool last_event_was_null) {
    const V3 mis_throughput = mis_weight1<S>(P);
    const float q = fminf(hmax(mis_throughput) * sqr(eta), .95f);
    const bool perform_rr = !last_event_was_null && depth > (uint32_t) sc.rr_depth;
    bool active = true;
    { const float u = smp.next_1d(); active &= !(u >= q && perform_rr); }
    update_weights<S>(P, mk3(q), mk3(1.f), channel, perform_rr);
    active &= depth < sc.max_depth;
    active &= any_nonzero(mis_weight1<S>(P));
    return active;
}

// the emitter sample of sample_emitter (volpathmis.cpp:474-490) + the weight it would have if nothing lies on the segment (TRACE_S adds it
// at once when the shadow ray starts outside any medium and hits nothing).  Writes the shadow ray and the record; false: no shadow ray.
template <bool S>
LR_DEV bool mis_nee_setup(const DScene &sc, const WVArgs &a, uint32_t slot, const MisW<S> &P, uint32_t channel, V3 ref_p, V3 ref_n, const SurfaceInteraction *ref_si,
                          int medium, const DirectionSample &ds, V3 weight, V3 val, float pdf_dir) {
    const V3 emitter_val = weight * ds.pdf;
    const Ray ray = spawn_ray_to(ref_p, ref_n, ds.p);
    if (ref_si && mesh_is_transition(sc, ref_si->mesh)) medium = target_medium(sc, *ref_si, ray.d);
    MisW<S> nee = P, uni = P;
    update_weights<S>(nee, mk3(ds.pdf), mk3(1.f), channel, true);
    update_weights<S>(nee, mk3(1.f), val, channel, true);
    update_weights<S>(uni, mk3(pdf_dir), val, channel, true);
    const V3 w = mis_weight2<S>(nee, uni) * emitter_val;
    float4 *m = mis_rec(a, slot);
    m[6] = make_float4(emitter_val.x, emitter_val.y, emitter_val.z, ds.pdf);
    m[7] = make_float4(val.x, val.y, val.z, pdf_dir);
    a.p.no(slot) = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.maxt);
    a.p.nd(slot) = make_float4(ray.d.x, ray.d.y, ray.d.z, ds.dist);
    a.p.nw(slot) = make_float4(w.x, w.y, w.z, __int_as_float(medium));
    return true;
}

// ------------------------------------------------------------------------------------------
// TRACK (volpathmis.cpp:236-312): free flights against the global majorant until a real scatter (1), the boundary (2) or the end (0).
// A persistent "while-while" loop like k_wv_track_dda: a path meets ~28 tentative collisions on the C3 frame and the chains of one warp differ
// widely, so a lane whose item is finished parks; once LR_MIS_REFILL lanes are parked they are written out (converged: the queue pushes use
// warp ballots) and pull the next queue entries.  One loop turn = one tentative collision per walking lane; per-lane draw order is unchanged.
// ------------------------------------------------------------------------------------------
#ifndef LR_MIS_REFILL
#define LR_MIS_REFILL 12              // refill once at least this many lanes of the warp are parked
#endif
template <int C, bool TO_FILM, bool S>
__global__ void __launch_bounds__(256, LR_MIS_OCC) k_mis_track(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    const WVPlanes &p = a.p;
    const uint32_t n_items = a.q->n_track;
    const unsigned lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    int event = -2;                                                        // -2 idle, -1 walking, 0 end of path, 1 scatter, 2 boundary
    uint32_t slot = 0, depth = 0, f = 0; float hit_t = 0.f, eta = 1.f; bool first = true, exhausted = false;
    Sampler smp; smp.state = smp.inc = 0; Ray ray; ray.o = ray.d = mk3(0.f); ray.maxt = 0.f;
    Hit h0; h0.t = LR_INF; h0.prim = 0xffffffffu; h0.u = h0.v = 0.f;
    MisW<S> P = misw_ones<S>(), N = misw_ones<S>();
    for (;;) {
        const unsigned walking = __ballot_sync(0xffffffffu, event == -1);
        if (__popc(walking) <= 32 - LR_MIS_REFILL) {
            if (__any_sync(0xffffffffu, event >= 0)) {
                if (event > 0) { float4 *mr = mis_rec(a, slot); misw_store<S>(mr, 0, P); misw_store<S>(mr, 3, N); }
                wv_track_out<C, TO_FILM>(sc, a, slot, event >= 0 ? event : -1, smp, ray, mk3(0.f), depth, hit_t, f, h0);
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
                    depth = __float_as_uint(p.thr(slot).w);
                    h0 = wf_hit(p.hit(slot)); hit_t = h0.t; eta = p.res(slot).w;
                    const float4 *mr = mis_rec(a, slot); P = misw_load<S>(mr, 0); N = misw_load<S>(mr, 3);
                    first = true; event = -1;
                }
            }
        }
        if (event == -1) {                                                  // one loop iteration of the reference for this lane
            const uint32_t channel = f & WF_CHANNEL_MASK; const int medium = wf_medium(f);
            const DMedium &m = sc.media[medium];
            if (!first && !mis_next_iteration<S>(sc, smp, P, eta, depth, channel, true)) event = 0;       // after a null collision: no roulette
            else {
                first = false;
                MediumInteraction mei = sample_interaction<false>(sc, medium, ray, smp.next_1d(), channel, nullptr, cnt.collisions);
                if (hit_t < mei.t) mei.t = LR_INF;
                const bool is_spectral = m.m.has_spectral_extinction != 0;
                if (is_spectral) {                                                  // :266-270
                    const float t = fminf(mei.t, hit_t) - mei.mint;
                    const V3 tr = exp3(mei.combined_extinction * (-t));
                    const V3 pdf = hit_t < mei.t ? tr : tr * mei.combined_extinction;
                    update_weights<S>(P, pdf, tr, channel, true);
                    update_weights<S>(N, pdf, tr, channel, true);
                }
                if (!mei_valid(mei)) event = 2;
                else {
                    const float null_scatter_prob = hmean(mei.sigma_n / mei.combined_extinction);
                    if (smp.next_1d() < null_scatter_prob) {                        // null collision: the reference's next loop iteration
                        if (is_spectral) {
                            update_weights<S>(P, mk3(null_scatter_prob), mei.sigma_n, channel, true);
                            update_weights<S>(N, mk3(1.f), mei.sigma_n, channel, true);
                        } else {
                            update_weights<S>(P, mei.sigma_n, mei.sigma_n, channel, true);
                            update_weights<S>(N, mk3(1.f), mk3(null_scatter_prob), channel, true);
                        }
                        ray.o = mei.p; hit_t = hit_t - mei.t;
                    } else {
                        depth += 1; cnt.real_scatters++; p.lsp(slot) = make_float4(mei.p.x, mei.p.y, mei.p.z, 0.f);
                        if (!(depth < sc.max_depth)) event = 0;
                        else {
                            if (is_spectral) update_weights<S>(P, mk3(1.f - null_scatter_prob), mei.sigma_s, channel, true);
                            else update_weights<S>(P, mei.sigma_t, mei.sigma_s, channel, true);
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

// medium scatter: emitter sampling (volpathmis.cpp:314-329)
template <bool S>
__global__ void __launch_bounds__(256, LR_MIS_OCC) k_mis_shade_a_medium(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a) {
    const WVPlanes &p = a.p;
    WF_FOR_QUEUE(a.q->n_scatter) {
        const bool item = qi < _n; bool nee = false; const uint32_t slot = item ? a.q_scatter[qi] : 0;
        if (item) {
            const float4 d4 = p.d(slot); uint32_t f = __float_as_uint(d4.w); const int medium = wf_medium(f);
            const DMedium &m = sc.media[medium]; Phase phase; phase.kind = m.m.phase; phase.g = m.m.phase_g;
            const bool sample_emitters = m.m.sample_emitters != 0; const uint32_t f0 = f;
            f |= WF_VALID_RAY; f = sample_emitters ? (f & ~WF_SPEC_CHAIN) : (f | WF_SPEC_CHAIN);
            if (sample_emitters) {
                Sampler smp = wv_rng(p, slot);
                const float4 o4 = p.o(slot); const V3 pt = mk3(o4.x, o4.y, o4.z), wi = -mk3(d4.x, d4.y, d4.z);
                V3 weight; const DirectionSample ds = sample_emitter_direction(sc, pt, smp.next_2d(), weight);
                wv_rng_store(p, slot, smp);
                if (ds.pdf != 0.f) {
                    const MisW<S> P = misw_load<S>(mis_rec(a, slot), 0);
                    const float phase_val = phase.eval_pdf(wi, ds.d);
                    nee = mis_nee_setup<S>(sc, a, slot, P, f & WF_CHANNEL_MASK, pt, mk3(0.f), nullptr, medium, ds, weight, mk3(phase_val), ds.delta ? 0.f : phase_val);
                }
            }
            if (f != f0) reinterpret_cast<float *>(&p.d(slot))[3] = __uint_as_float(f);
        }
        wf_push(a.q_shadow, &a.q->n_shadow, slot, nee);
    }
}

// surface (volpathmis.cpp:352-416); every item has a valid hit
template <int C, bool TO_FILM, bool S>
__global__ void __launch_bounds__(128, 2 * LR_MIS_OCC) k_mis_shade_a_surface(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a, int bucket) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    const WVPlanes &p = a.p;
    WF_FOR_QUEUE(a.q->n_surf[bucket]) {
        const bool item = qi < _n; const uint32_t slot = item ? a.q_surf[bucket][qi] : 0; bool nee = false, gone = false; uint32_t f = 0;
        if (item) {
            const float4 o4 = p.o(slot), d4 = p.d(slot);
            f = __float_as_uint(d4.w); const uint32_t depth = __float_as_uint(p.thr(slot).w); const int medium = wf_medium(f); const uint32_t channel = f & WF_CHANNEL_MASK;
            Ray ray; ray.o = mk3(o4.x, o4.y, o4.z); ray.d = mk3(d4.x, d4.y, d4.z); ray.maxt = o4.w;
            Hit h = wf_hit(p.hit(slot));
            SurfaceInteraction si = compute_si(sc, ray, h, sc.tris_gid);
            if (sc.hide_emitters && medium < 0 && depth == 0u && sc.meshes[si.mesh].emitter >= 0) {                     // :360-375
                Ray r2 = spawn_ray(si.p, si.n, ray.d); Hit pi = skip_area_emitters(sc, sc.tris_gid, r2, cnt);
                si = compute_si(sc, r2, pi, sc.tris_gid); h = pi; p.hit(slot) = wf_hit4(h);
                p.o(slot) = make_float4(r2.o.x, r2.o.y, r2.o.z, r2.maxt);
            }
            float4 *mr = mis_rec(a, slot);
            const MisW<S> P = misw_load<S>(mr, 0);
            const bool count_direct = depth == 0u || (f & WF_SPEC_CHAIN);
            const int emitter = emitter_of(sc, si);
            if (emitter >= 0 && !(depth == 0u && sc.hide_emitters)) {                                                     // :377-398
                MisW<S> N = misw_load<S>(mr, 3);
                const float4 l4 = p.lsp(slot);
                const DirectionSample ds = ds_from_si(sc, si, mk3(l4.x, l4.y, l4.z));
                update_weights<S>(N, mk3(pdf_emitter_direction(sc, ds)), mk3(1.f), channel, true);      // masked by active_e only (:387-392)
                misw_store<S>(mr, 3, N);
                const V3 emitted = eval_emitter_at(sc, emitter, si);
                const V3 c = count_direct ? mis_weight1<S>(P) * emitted : mis_weight2<S>(P, N) * emitted;
                float4 r4 = p.res(slot); r4.x += c.x; r4.y += c.y; r4.z += c.z; p.res(slot) = r4;
            }
            if (!si_valid(si)) {
                gone = true; f |= WF_DEAD; reinterpret_cast<float *>(&p.d(slot))[3] = __uint_as_float(f);
            } else {
                BSDF bsdf(sc, si);
                if ((bsdf.flags() & F_SMOOTH) && (depth + 1 < sc.max_depth)) {                                           // :403-416
                    Sampler smp = wv_rng(p, slot);
                    V3 weight; const DirectionSample ds = sample_emitter_direction(sc, si.p, smp.next_2d(), weight);
                    wv_rng_store(p, slot, smp);
                    if (ds.pdf != 0.f) {
                        const V3 wo = si.sh.to_local(ds.d); float bsdf_pdf; const V3 bsdf_val = bsdf.eval_pdf(si.wi, wo, bsdf_pdf);
                        nee = mis_nee_setup<S>(sc, a, slot, P, channel, si.p, si.n, &si, medium, ds, weight, bsdf_val, ds.delta ? 0.f : bsdf_pdf);
                    }
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

// rays that left the scene (volpathmis.cpp:377-398 with an invalid si)
template <int C, bool TO_FILM, bool S>
__global__ void __launch_bounds__(256, LR_MIS_OCC) k_mis_escape(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a) {
    const WVPlanes &p = a.p;
    WF_FOR_QUEUE(a.q->n_escape) {
        const bool item = qi < _n; const uint32_t slot = item ? a.q_escape[qi] : 0;
        V3 L = mk3(0.f); uint32_t f = 0; float4 lp = make_float4(0.f, 0.f, 0.f, 0.f);
        if (item) {
            const float4 d4 = p.d(slot), r4 = p.res(slot); lp = p.lane(slot);
            f = __float_as_uint(d4.w); const uint32_t depth = __float_as_uint(p.thr(slot).w);
            L = mk3(r4.x, r4.y, r4.z);
            if (sc.env >= 0 && !(depth == 0u && sc.hide_emitters)) {
                const V3 dir = mk3(d4.x, d4.y, d4.z);
                const bool count_direct = depth == 0u || (f & WF_SPEC_CHAIN);
                const float4 *mr = mis_rec(a, slot);
                const MisW<S> P = misw_load<S>(mr, 0); MisW<S> N = misw_load<S>(mr, 3);
                DirectionSample ds; ds.d = dir; ds.emitter = sc.env; ds.dist = 0.f; ds.n = ds.p = mk3(0.f); ds.uv = mk2(0.f, 0.f); ds.pdf = 0.f; ds.delta = false;
                update_weights<S>(N, mk3(pdf_emitter_direction(sc, ds)), mk3(1.f), f & WF_CHANNEL_MASK, true);
                const V3 emitted = emitter_eval(sc, sc.env, -dir.z, dir);
                L += count_direct ? mis_weight1<S>(P) * emitted : mis_weight2<S>(P, N) * emitted;
            }
        }
        wv_finish<C, TO_FILM>(sc, a, item, lp, L, f);
        wf_push(a.q_regen, &a.q->n_regen, slot, item && wv_more_passes(a, f));
    }
}

// NEE: the walk of sample_emitter (volpathmis.cpp:523-629) from TRACE_S's closest hit, then the two end-point updates and the estimate.
// Persistent refill loop like k_mis_track: one loop turn = one iteration of the reference's walk per lane.
template <bool S>
__global__ void __launch_bounds__(256, LR_MIS_OCC) k_mis_nee(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    const WVPlanes &p = a.p;
    const uint32_t n_items = a.q->n_nee;
    const unsigned lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    int state = -2;                                                        // -2 idle, -1 walking, 0 walk finished
    uint32_t slot = 0, channel = 0; bool exhausted = false, needs_intersection = false;
    Sampler smp; smp.state = smp.inc = 0; Ray ray; ray.o = ray.d = mk3(0.f); ray.maxt = 0.f;
    float max_dist = 0.f, ds_dist = 0.f, total_dist = 0.f, si_t = LR_INF; int medium = -1;
    Hit h; h.t = LR_INF; h.prim = 0xffffffffu; h.u = h.v = 0.f;
    MisW<S> nee = misw_ones<S>(), uni = misw_ones<S>();
    for (;;) {
        const unsigned walking = __ballot_sync(0xffffffffu, state == -1);
        if (__popc(walking) <= 32 - LR_MIS_REFILL) {
            if (state == 0) {                                               // the two end-point updates and the estimate (:326-329, :413-416)
                const float4 *mr = mis_rec(a, slot); const float4 e4 = mr[6], v4 = mr[7];
                const V3 val = mk3(v4.x, v4.y, v4.z);
                update_weights<S>(nee, mk3(1.f), val, channel, true);
                update_weights<S>(uni, mk3(v4.w), val, channel, true);
                const V3 c = mis_weight2<S>(nee, uni) * mk3(e4.x, e4.y, e4.z);
                float4 r4 = p.res(slot); r4.x += c.x; r4.y += c.y; r4.z += c.z; p.res(slot) = r4;
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
                    h = wf_hit(p.nhit(slot)); si_t = h.t; needs_intersection = false;
                    const float4 *mr = mis_rec(a, slot);
                    nee = misw_load<S>(mr, 0); uni = nee;
                    update_weights<S>(nee, mk3(mr[6].w), mk3(1.f), channel, true);     // :481
                    state = -1;
                }
            }
        }
        if (state == -1) {                                                  // one iteration of the walk
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
                    bool is_spectral = m.m.has_spectral_extinction != 0, not_spectral = !is_spectral;
                    if (is_spectral) {
                        const float t = fminf(remaining_dist, fminf(mei.t, si_t)) - mei.mint;
                        const V3 tr = exp3(mei.combined_extinction * (-t));
                        const V3 ffp = (si_t < mei.t || mei.t > remaining_dist) ? tr : tr * mei.combined_extinction;
                        update_weights<S>(nee, ffp, tr, channel, true);
                        update_weights<S>(uni, ffp, tr, channel, true);
                    }
                    if (mei.t > remaining_dist && mei_valid(mei)) total_dist = ds_dist;
                    if (mei.t > remaining_dist) mei.t = LR_INF;
                    escaped_medium = !mei_valid(mei); active_medium = mei_valid(mei);
                    is_spectral &= active_medium; not_spectral &= active_medium;
                    if (active_medium) {
                        total_dist += mei.t; ray.o = mei.p; si_t = si_t - mei.t;
                        if (is_spectral) {
                            update_weights<S>(nee, mk3(1.f), mei.sigma_n, channel, true);
                            update_weights<S>(uni, mk3(hmean(mei.sigma_n / mei.combined_extinction)), mei.sigma_n, channel, true);
                        }
                        if (not_spectral) {
                            update_weights<S>(nee, mk3(1.f), mei.sigma_n / mei.combined_extinction, channel, true);
                            update_weights<S>(uni, mei.sigma_n, mei.sigma_n, channel, true);
                        }
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
                    const V3 bsdf_val = mk3(bsdf.eval_null_transmission());
                    update_weights<S>(nee, mk3(1.f), bsdf_val, channel, true);
                    update_weights<S>(uni, mk3(1.f), bsdf_val, channel, true);
                    Ray r2 = spawn_ray(si.p, si.n, ray.d); ray.o = r2.o;
                }
                needs_intersection |= active_surface;
                if (S) active &= (active_medium || active_surface) && any_nonzero(mis_weight1<S>(uni));
                else active &= (active_medium || active_surface) && (any_nonzero(uni.r[0]) || any_nonzero(nee.r[0]));
                if (active_surface && mesh_is_transition(sc, si.mesh)) medium = target_medium(sc, si, ray.d);
            }
            if (!active) state = 0;
        }
        if (exhausted && !__any_sync(0xffffffffu, state != -2)) break;
    }
    flush_counters(cnt, 0, a.stats);
}

// medium scatter: phase-function sampling (volpathmis.cpp:331-345), then the top of the next iteration
template <int C, bool TO_FILM, bool S>
__global__ void __launch_bounds__(256, LR_MIS_OCC) k_mis_shade_b_medium(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a) {
    const WVPlanes &p = a.p;
    WF_FOR_QUEUE(a.q->n_scatter) {
        const bool item = qi < _n; const uint32_t slot = item ? a.q_scatter[qi] : 0;
        bool alive = false; uint32_t f = 0; V3 L = mk3(0.f); float4 lp = make_float4(0.f, 0.f, 0.f, 0.f);
        if (item) {
            Sampler smp = wv_rng(p, slot);
            const float4 d4 = p.d(slot); f = __float_as_uint(d4.w); const uint32_t channel = f & WF_CHANNEL_MASK;
            const DMedium &m = sc.media[wf_medium(f)]; Phase phase; phase.kind = m.m.phase; phase.g = m.m.phase_g;
            float4 *mr = mis_rec(a, slot);
            MisW<S> P = misw_load<S>(mr, 0), N = P;                                    // p_over_f_nee = p_over_f (:331)
            (void) smp.next_1d(); const V2 s2 = smp.next_2d();
            float phase_pdf; const V3 wo = phase.sample(-mk3(d4.x, d4.y, d4.z), s2, phase_pdf);
            update_weights<S>(P, mk3(phase_pdf), mk3(1.f * phase_pdf), channel, true);
            update_weights<S>(N, mk3(1.f), mk3(1.f * phase_pdf), channel, true);
            const uint32_t depth = __float_as_uint(p.thr(slot).w);
            const float4 r4 = p.res(slot);
            alive = mis_next_iteration<S>(sc, smp, P, r4.w, depth, channel, false);
            wv_rng_store(p, slot, smp);
            if (alive) {
                p.d(slot) = make_float4(wo.x, wo.y, wo.z, d4.w);
                reinterpret_cast<float *>(&p.o(slot))[3] = LR_LARGEST;
                misw_store<S>(mr, 0, P); misw_store<S>(mr, 3, N);
            } else { L = mk3(r4.x, r4.y, r4.z); lp = p.lane(slot); }
        }
        wv_finish<C, TO_FILM>(sc, a, item && !alive, lp, L, f);
        wf_push(a.q_trace, &a.q->n_trace, slot, alive);
        wf_push(a.q_regen, &a.q->n_regen, slot, item && !alive && wv_more_passes(a, f));
    }
}

// surface: BSDF sampling, next ray, medium transition (volpathmis.cpp:418-446), then the top of the next iteration
template <int C, bool TO_FILM, bool S>
__global__ void __launch_bounds__(128, 2 * LR_MIS_OCC) k_mis_shade_b_surface(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a, int bucket) {
    const WVPlanes &p = a.p;
    WF_FOR_QUEUE(a.q->n_surf[bucket]) {
        bool item = qi < _n; const uint32_t slot = item ? a.q_surf[bucket][qi] : 0;
        bool alive = false; uint32_t f = 0; V3 L = mk3(0.f); float4 lp = make_float4(0.f, 0.f, 0.f, 0.f);
        float4 d4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (item) { d4 = p.d(slot); f = __float_as_uint(d4.w); item = !(f & WF_DEAD); }
        if (item) {
            Sampler smp = wv_rng(p, slot);
            const float4 o4 = p.o(slot), r4 = p.res(slot); const uint32_t channel = f & WF_CHANNEL_MASK;
            Ray ray; ray.o = mk3(o4.x, o4.y, o4.z); ray.d = mk3(d4.x, d4.y, d4.z); ray.maxt = o4.w;
            const SurfaceInteraction si = compute_si(sc, ray, wf_hit(p.hit(slot)), sc.tris_gid);
            BSDF bsdf(sc, si);
            const float s1 = smp.next_1d(); const V2 s2 = smp.next_2d();
            V3 bsdf_weight; const BSDFSample bs = bsdf.sample(si.wi, s1, s2, bsdf_weight);
            const bool invalid_bsdf_sample = bs.pdf == 0.f, active_surface = bs.pdf > 0.f;
            float eta = r4.w; uint32_t depth = __float_as_uint(p.thr(slot).w);
            const bool non_null = active_surface && !(bs.sampled_type & F_NULL);
            if (non_null || invalid_bsdf_sample) f |= WF_VALID_RAY;
            bool spec = (f & WF_SPEC_CHAIN) != 0;
            spec |= non_null && (bs.sampled_type & F_DELTA); spec &= !(active_surface && (bs.sampled_type & F_SMOOTH));
            f = spec ? (f | WF_SPEC_CHAIN) : (f & ~WF_SPEC_CHAIN);
            float4 *mr = mis_rec(a, slot);
            MisW<S> P = misw_load<S>(mr, 0), N = misw_load<S>(mr, 3);
            if (non_null) { depth += 1u; p.lsp(slot) = make_float4(si.p.x, si.p.y, si.p.z, 0.f); N = P; }
            update_weights<S>(P, mk3(bs.pdf), bsdf_weight * bs.pdf, channel, active_surface);
            update_weights<S>(N, mk3(1.f), bsdf_weight * bs.pdf, channel, non_null);
            if (active_surface) {
                eta *= bs.eta;
                ray = spawn_ray(si.p, si.n, si.sh.to_world(bs.wo));
                if (mesh_is_transition(sc, si.mesh)) f = wf_set_medium(f, target_medium(sc, si, ray.d));
                alive = mis_next_iteration<S>(sc, smp, P, eta, depth, channel, false);
            }
            wv_rng_store(p, slot, smp);
            if (alive) {
                p.o(slot) = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.maxt);
                p.d(slot) = make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float(f));
                reinterpret_cast<float *>(&p.thr(slot))[3] = __uint_as_float(depth);
                if (eta != r4.w) reinterpret_cast<float *>(&p.res(slot))[3] = eta;
                misw_store<S>(mr, 0, P); misw_store<S>(mr, 3, N);
            } else { L = mk3(r4.x, r4.y, r4.z); lp = p.lane(slot); }
        }
        wv_finish<C, TO_FILM>(sc, a, item && !alive, lp, L, f);
        wf_push(a.q_trace, &a.q->n_trace, slot, alive);
        wf_push(a.q_regen, &a.q->n_regen, slot, item && !alive && wv_more_passes(a, f));
    }
}

} // namespace lr
