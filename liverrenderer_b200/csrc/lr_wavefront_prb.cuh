// lr_wavefront_prb.cuh -- wavefront `prbvolpath`: sample_prb<PRIMAL>() / sample_emitter_prb<ADJ>() of lr_prb.cuh
// (reference: src/python/python/ad/integrators/prbvolpath.py:96-444) cut into the same stages and queues as the
// wavefront volpath (lr_wavefront.cuh).  ADJ = false: primal pass (radiance per lane -> film or per-lane buffer).
// ADJ = true: adjoint replay -- the slot's radiance starts at the lane's primal result and every contribution is
// subtracted as it is replayed; each collision scatters
//     d loss / d grid += w_k * sum_c dL_c * (L_c / weight_c) * d weight_c / d sigma_t * scale
// into the gradient grid with the warp-aggregated footprint scatter (grid_grad_scatter, lr_prb.cuh).
// Both passes draw the lane's PCG32 stream in the reference's order, so wavefront == per-lane kernel == oracle.
#pragma once
#include "lr_wavefront.cuh"

#ifndef LR_WFP_OCC
#define LR_WFP_OCC 4
#endif

namespace lr {

// The detached-sampling gradient of a null collision is dL * L * (-scale) / sigma_n: singular where the majorant is tight.
// The PRB passes therefore walk the super-grid with slackened local majorants (sigma_n >= mu / 2); primal `volpath` keeps
// the tight ones.  Both PRB passes use the same factor, so the adjoint replays exactly the primal paths.
static constexpr float kPrbMajorantSlack = 2.f;

struct WFPrbArgs {                     // extra inputs of the PRB passes
    const float *grad_in; const float *weight_img; uint32_t grad_channels;      // adjoint: dLoss/dImage, per-pixel filter-weight sums
    const float *primal_L;            // adjoint: float4 per lane (lane - lane_begin), written by the primal pass
};

// ------------------------------------------------------------------------------------------
// GEN (common.py:231-309 sample_rays; adjoint: delta L of the lane = adjoint of splat + develop, common.py:370-400)
// ------------------------------------------------------------------------------------------
template <bool ADJ>
__global__ void __launch_bounds__(256, LR_WFP_OCC) k_wfp_gen(const __grid_constant__ DScene sc, const __grid_constant__ WFArgs a, const __grid_constant__ WFPrbArgs pa) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0; unsigned long long nsamples = 0;
    uint32_t *q_out = a.q_active[a.cur ^ 1], *n_out = &a.q->n_active[a.cur ^ 1];
    const bool box = sc.sensor.rfilter == LR_RFILTER_BOX, coalesce = a.spp_per_pass >= 4;
    WF_FOR_QUEUE(a.q->n_regen) {
        bool item = qi < _n; uint32_t slot = item ? a.q_regen[qi] : 0; bool live = false; uint64_t lane = 0, li = 0;
        WState *S = a.st + slot;
        {
            unsigned m = __ballot_sync(0xffffffffu, item); unsigned l = threadIdx.x & 31; unsigned long long base = 0;
            if (m) {
                if (l == (unsigned) (__ffs(m) - 1)) base = atomicAdd(&a.q->lane_next, (unsigned long long) __popc(m));
                base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
                if (item) { li = a.lane_begin + base + __popc(m & ((1u << l) - 1u)); live = li < a.lane_end; lane = lane_global(a.map, li); }
            }
        }
        if (item && !live) S->lsp_flags.w = __uint_as_float(WF_DEAD);
        if (live) {
            LaneSetup L; seed_sampler(sc, L.smp, a.seed, lane, a.spp_per_pass, 1);
            lane_begin_sample(sc, L, lane, a.spp_per_pass); lane_next_ray(sc, L);
            V2 fp2 = box ? L.pos : L.sample_pos;
            uint32_t channel = (uint32_t) fminf(3.f * L.smp.next_1d(), 2.f);                 // prbvolpath.py:136-139
            uint32_t f = channel | WF_HAS_LANE | WF_SPEC_CHAIN;                               // valid_ray = false, specular_chain = true
            f = wf_set_medium(f, -1);                                                         // :127-128: the sensor is never inside a medium
            cnt.rays++; Hit h = bvh_trace<false, false>(sc, L.ray); nsamples++;
            V3 Lin = mk3(0.f), dL = mk3(0.f);
            if (ADJ) {
                float4 p = reinterpret_cast<const float4 *>(pa.primal_L)[li - a.lane_begin]; Lin = mk3(p.x, p.y, p.z);
                Footprint fp; const int W = (int) sc.sensor.film_width, H = (int) sc.sensor.film_height;
                if (film_footprint(sc, fp2, coalesce, fp)) {
                    for (int j = 0; j < fp.ny; ++j) { int y = fp.y0 + j; if (y < 0 || y >= H) continue;
                        for (int i = 0; i < fp.nx; ++i) { int x = fp.x0 + i; if (x < 0 || x >= W) continue;
                            float w = __fmul_rn(fp.wy[j], fp.wx[i]); float ws = pa.weight_img[(size_t) y * W + x]; float d = ws == 0.f ? 1.f : ws;
                            const float *g = pa.grad_in + ((size_t) y * W + x) * pa.grad_channels;
                            float s = w / d; dL.x = fmaf(g[0], s, dL.x); dL.y = fmaf(g[1], s, dL.y); dL.z = fmaf(g[2], s, dL.z); } }
                }
            }
            S->rng_state = L.smp.state; S->rng_inc = L.smp.inc;
            S->o_maxt = make_float4(L.ray.o.x, L.ray.o.y, L.ray.o.z, L.ray.maxt); S->d_eta = make_float4(L.ray.d.x, L.ray.d.y, L.ray.d.z, 1.f);
            S->thr_lpdf = make_float4(1.f, 1.f, 1.f, 1.f); S->res_depth = make_float4(Lin.x, Lin.y, Lin.z, __uint_as_float(0u));
            S->hit = wf_hit4(h); S->lsp_flags = make_float4(0.f, 0.f, 0.f, __uint_as_float(f));
            S->lane_pos = make_float4(__uint_as_float((uint32_t) lane), __uint_as_float((uint32_t) (lane >> 32)), fp2.x, fp2.y);
            S->pad0 = make_float4(dL.x, dL.y, dL.z, 0.f);
        }
        wf_push(q_out, n_out, slot, live);
    }
    flush_counters(cnt, nsamples, a.stats);
}

// adjoint of one tentative collision of the main walk (prbvolpath.py:169-204 + the backward of `weight`)
LR_DEV void wfp_scatter_main(const DScene &sc, const DMedium &m, const GridTap &tap, V3 dL, V3 L, V3 weight, V3 dweight) {
    const DGrid &g = sc.grids[m.m.sigma_t_grid];
    float gsum = (dL.x * (L.x / fmaxf(1e-8f, weight.x))) * dweight.x + (dL.y * (L.y / fmaxf(1e-8f, weight.y))) * dweight.y +
                 (dL.z * (L.z / fmaxf(1e-8f, weight.z))) * dweight.z;
    if (g.differentiable && g.ch == 1 && gsum != 0.f) grid_grad_scatter(g, tap, gsum, true);
}

// ------------------------------------------------------------------------------------------
// TRACK (prbvolpath.py:141-204)
// ------------------------------------------------------------------------------------------
template <int C, bool TO_FILM, bool ADJ>
__global__ void __launch_bounds__(256, LR_WFP_OCC) k_wfp_track(const __grid_constant__ DScene sc, const __grid_constant__ WFArgs a) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    const uint32_t *q_in = a.q_active[a.cur];
    WF_FOR_QUEUE(a.q->n_active[a.cur]) {
        bool item = qi < _n; uint32_t slot = item ? q_in[qi] : 0; WState *S = a.st + slot;
        int event = -1;
        if (item) {
            Sampler smp; smp.state = S->rng_state; smp.inc = S->rng_inc;
            float4 t4 = S->o_maxt; Ray ray; ray.o = mk3(t4.x, t4.y, t4.z); ray.maxt = t4.w;
            t4 = S->d_eta; ray.d = mk3(t4.x, t4.y, t4.z); float eta = t4.w;
            t4 = S->thr_lpdf; V3 thr = mk3(t4.x, t4.y, t4.z); float lpdf = t4.w;
            float4 h4 = S->hit; float hit_t = h4.x;
            float4 lf = S->lsp_flags; uint32_t f = __float_as_uint(lf.w); V3 lsp = mk3(lf.x, lf.y, lf.z);
            float4 r4 = S->res_depth; uint32_t depth = __float_as_uint(r4.w); const V3 Lcur = mk3(r4.x, r4.y, r4.z);
            V3 dL = mk3(0.f); if (ADJ) { float4 d4 = S->pad0; dL = mk3(d4.x, d4.y, d4.z); }
            const uint32_t channel = f & WF_CHANNEL_MASK; const int medium = wf_medium(f);
            for (;;) {
                bool active = any_nonzero(thr);
                float q = fminf(hmax(thr) * sqr(eta), 0.99f);
                bool perform_rr = depth > (uint32_t) sc.rr_depth;
                if (active) { float u = smp.next_1d(); active &= (u < q) || !perform_rr; }
                if (perform_rr) thr *= 1.f / q;
                if (!active) { event = 0; break; }
                if (medium < 0) { event = 2; break; }
                const DMedium &m = sc.media[medium]; const bool hetero = m.m.kind == LR_MEDIUM_HETEROGENEOUS;
                GridTap tap; V3 weight, dweight = mk3(0.f); bool valid, null_c = false; V3 p = mk3(0.f); float t_c = 0.f;
                V3 sigma_t, sigma_n, ce; float inv_pdf_c = 1.f;            // inv_pdf_c: Tr/pdf of the free flight (channel-independent here)
                if (medium_uses_dda(sc, m)) {
                    // local-majorant walk up to the next tentative collision (or the end of the segment)
                    const DGrid &g = sc.grids[m.m.sigma_t_grid]; float mint, maxt; valid = false; float mu = 0.f;
                    if (medium_clip(g, ray, mint, maxt)) {
                        const float t_end = fminf(maxt, hit_t);
                        if (mint < t_end) {
                            DDA dd; dda_init(g, ray, mint, dd); float t = mint, tau = -logf(1.f - smp.next_1d());
                            valid = dda_walk(g, m.m.scale * kPrbMajorantSlack, dd, t, tau, t_end, mu); t_c = t;
                        }
                    }
                    if (!valid) { event = 2; break; }
                    cnt.collisions++; p = ray_at(ray, t_c);
                    sigma_t = grid_eval<ADJ>(g, p, &tap) * m.m.scale; ce = mk3(mu); sigma_n = ce - sigma_t; inv_pdf_c = 1.f / mu;
                    weight = mk3(inv_pdf_c);
                } else {
                    float u = smp.next_1d();
                    MediumInteraction mei = sample_interaction<ADJ>(sc, medium, ray, u, channel, &tap, cnt.collisions);
                    if (hit_t < mei.t) mei.t = LR_INF;
                    float t = fminf(mei.t, hit_t) - mei.mint;
                    V3 tr = exp3(mei.combined_extinction * (-t));
                    V3 ffp = hit_t < mei.t ? tr : tr * mei.combined_extinction;
                    float tr_pdf = idx3(ffp, channel);
                    weight = tr_pdf > 0.f ? tr / tr_pdf : mk3(0.f);
                    valid = mei_valid(mei);
                    if (!valid) { thr *= weight; event = 2; break; }
                    p = mei.p; t_c = mei.t; sigma_t = mei.sigma_t; sigma_n = mei.sigma_n; ce = mei.combined_extinction;
                }
                // a tentative collision at p: null or real (prbvolpath.py:181-204)
                float scatter_prob = 1.f;
                if (sc.prb_handle_null) { scatter_prob = hmean(sigma_t / ce); null_c = smp.next_1d() >= scatter_prob; }
                if (null_c) {
                    V3 A = weight; weight *= sigma_n / (1.f - scatter_prob); dweight = A * (-m.m.scale / (1.f - scatter_prob));
                    thr *= weight;
                    if (ADJ && hetero) wfp_scatter_main(sc, m, tap, dL, Lcur, weight, dweight);
                    ray.o = p; hit_t = hit_t - t_c;
                    if (!(depth < sc.max_depth)) { event = 0; break; }      // :189-190: `active &= depth < max_depth` ends the lane after this collision
                    continue;
                }
                depth += 1; lsp = p; cnt.real_scatters++;
                if (!(depth < sc.max_depth)) { thr *= weight; event = 0; break; }
                {
                    V3 A = weight; V3 albedo = medium_albedo(sc, m, p);
                    weight *= (sigma_t * albedo) / scatter_prob; dweight = A * albedo * (m.m.scale / scatter_prob);
                    thr *= weight;
                    if (ADJ && hetero) wfp_scatter_main(sc, m, tap, dL, Lcur, weight, dweight);
                }
                ray.o = p; hit_t = hit_t - t_c; event = 1; break;
            }
            S->rng_state = smp.state; S->rng_inc = smp.inc;
            if (event == 0) {
                if (!ADJ) wf_finish<C, TO_FILM>(sc, a, *S, Lcur, f);
            } else {
                if (event == 2) event = 2 + wf_surf_bucket(sc, a.tris_gid, wf_hit(h4));
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

// NEE set-up (prbvolpath.py:354-378): emitter direction sample, walk medium, shadow-ray closest hit
LR_DEV bool wfp_nee_setup(const DScene &sc, const WFArgs &a, WState *S, Sampler &smp, V3 ref_p, V3 ref_n, const SurfaceInteraction *ref_si,
                          int medium, DirectionSample &ds, Counters &cnt) {
    V3 emitter_val; ds = sample_emitter_direction(sc, ref_p, smp.next_2d(), emitter_val);
    if (ds.pdf == 0.f) return false;
    if (ref_si && mesh_is_transition(sc, ref_si->mesh)) medium = target_medium(sc, *ref_si, ds.d);
    Ray ray = spawn_ray_to(ref_p, ref_n, ds.p);
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

// SHADE_A, medium scatter (prbvolpath.py:259-291, medium part)
__global__ void __launch_bounds__(256, LR_WFP_OCC) k_wfp_shade_a_medium(const __grid_constant__ DScene sc, const __grid_constant__ WFArgs a) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    WF_FOR_QUEUE(a.q->n_scatter) {
        bool item = qi < _n, nee = false; uint32_t slot = item ? a.q_scatter[qi] : 0;
        if (item) {
            WState *S = a.st + slot;
            Sampler smp; smp.state = S->rng_state; smp.inc = S->rng_inc;
            uint32_t f = wf_flags(*S); int medium = wf_medium(f);
            const DMedium &m = sc.media[medium]; Phase phase; phase.kind = m.m.phase; phase.g = m.m.phase_g;
            bool sample_emitters = m.m.sample_emitters != 0;
            f = sample_emitters ? (f & ~WF_SPEC_CHAIN) : (f | WF_SPEC_CHAIN); f &= ~(WF_NEE_ACTIVE | WF_DEAD);
            if (sample_emitters) {
                float4 o = S->o_maxt, d = S->d_eta, t = S->thr_lpdf; V3 p = mk3(o.x, o.y, o.z), wi = -mk3(d.x, d.y, d.z), thr = mk3(t.x, t.y, t.z);
                DirectionSample ds;
                if (wfp_nee_setup(sc, a, S, smp, p, mk3(0.f), nullptr, medium, ds, cnt)) {
                    float pv = phase.eval_pdf(wi, ds.d); V3 w = thr * pv;
                    S->ntr_mis = make_float4(1.f, 1.f, 1.f, mis_weight(ds.pdf, ds.delta ? 0.f : pv));
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

// SHADE_A, surface / escape (prbvolpath.py:206-291, surface part).  sgn = +1 primal, -1 adjoint.
template <int C, bool TO_FILM, bool ADJ>
__global__ void __launch_bounds__(128, 2 * LR_WFP_OCC) k_wfp_shade_a_surface(const __grid_constant__ DScene sc, const __grid_constant__ WFArgs a, int bucket) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0; const float sgn = ADJ ? -1.f : 1.f;
    WF_FOR_QUEUE(a.q->n_surf[bucket]) {
        bool item = qi < _n; uint32_t slot = item ? a.q_surf[bucket][qi] : 0; bool dead = false, nee = false;
        if (item) {
            WState *S = a.st + slot;
            Sampler smp; smp.state = S->rng_state; smp.inc = S->rng_inc;
            uint32_t f = wf_flags(*S) & ~(WF_NEE_ACTIVE | WF_DEAD); int medium = wf_medium(f);
            float4 o = S->o_maxt, d = S->d_eta, t = S->thr_lpdf, r4 = S->res_depth;
            Ray ray; ray.o = mk3(o.x, o.y, o.z); ray.d = mk3(d.x, d.y, d.z); ray.maxt = o.w;
            V3 thr = mk3(t.x, t.y, t.z), L = mk3(r4.x, r4.y, r4.z); uint32_t depth = __float_as_uint(r4.w);
            Hit h = wf_hit(S->hit);
            SurfaceInteraction si = compute_si(sc, ray, h, a.tris_gid);
            if (sc.hide_emitters && medium < 0 && depth == 0u && si_valid(si) && sc.meshes[si.mesh].emitter >= 0) {        // :218-231
                Ray r2 = spawn_ray(si.p, si.n, ray.d); Hit pi = skip_area_emitters(sc, a.tris_gid, r2, cnt);
                si = compute_si(sc, ray, pi, a.tris_gid); h = pi; S->hit = wf_hit4(h);
            }
            bool count_direct = depth == 0u || (f & WF_SPEC_CHAIN);
            int emitter = emitter_of(sc, si);
            if (emitter >= 0 && !(depth == 0u && sc.hide_emitters)) {                                                     // :233-257
                float4 l = S->lsp_flags; DirectionSample ds = ds_from_si(sc, si, mk3(l.x, l.y, l.z));
                float emitter_pdf = pdf_emitter_direction(sc, ds);
                V3 emitted = eval_emitter_at(sc, emitter, si);
                V3 contrib = count_direct ? thr * emitted : thr * mis_weight(t.w, emitter_pdf) * emitted;
                L += contrib * sgn;
            }
            // a lane that arrives with depth == max_depth still counts the emitter it sees (`active_surface` is not re-masked
            // after `active &= depth < max_depth`, prbvolpath.py:189-257) and ends here
            if (!si_valid(si) || !(depth < sc.max_depth)) {
                dead = true; f |= WF_DEAD; if (!ADJ) wf_finish<C, TO_FILM>(sc, a, *S, L, f);
            } else {
                BSDF bsdf(sc, si);
                if ((bsdf.flags() & F_SMOOTH) && (depth + 1 < sc.max_depth)) {
                    DirectionSample ds;
                    if (wfp_nee_setup(sc, a, S, smp, si.p, si.n, &si, medium, ds, cnt)) {
                        float nee_pdf; V3 w = thr * bsdf.eval_pdf(si.wi, si.sh.to_local(ds.d), nee_pdf);
                        S->ntr_mis = make_float4(1.f, 1.f, 1.f, mis_weight(ds.pdf, ds.delta ? 0.f : nee_pdf));
                        S->nw_med.x = w.x; S->nw_med.y = w.y; S->nw_med.z = w.z;
                        f |= WF_NEE_ACTIVE;
                    }
                }
            }
            S->rng_state = smp.state; S->rng_inc = smp.inc;
            S->res_depth = make_float4(L.x, L.y, L.z, r4.w);
            S->lsp_flags.w = __uint_as_float(f); nee = (f & WF_NEE_ACTIVE) != 0;
        }
        wf_push(a.q_regen, &a.q->n_regen, slot, dead);
        wf_push(a.q_nee, &a.q->n_nee, slot, nee);
    }
    flush_counters(cnt, 0, a.stats);
}

// One shadow-segment walk of PRBVolpathIntegrator.sample_emitter (prbvolpath.py:379-444).  SCATTER = false: returns the
// transmittance.  SCATTER = true: replays the walk (same sampler state) and back-propagates `adj` = the NEE contribution
// through every tr_multiplier = sigma_n / majorant:  d/d grid = dL * adj / tr_mult * (-scale / majorant).
template <bool SCATTER>
LR_DEV V3 wfp_nee_walk(const DScene &sc, const WFArgs &a, const WState *S, Sampler &smp, uint32_t channel, V3 adj, V3 dL, Counters &cnt) {
    float4 o = S->no_maxd, d = S->nd_total; Ray ray; ray.o = mk3(o.x, o.y, o.z); ray.d = mk3(d.x, d.y, d.z);
    const float max_dist = o.w; float total_dist = 0.f;
    int medium = __float_as_int(S->nw_med.w);
    Hit h = wf_hit(S->nhit); float si_t = h.t; bool needs_intersection = false, active = true;
    V3 transmittance = mk3(1.f);
    while (active) {
        float remaining_dist = max_dist - total_dist; ray.maxt = remaining_dist;
        active &= remaining_dist > 0.f;
        if (!active) break;
        if (needs_intersection) { cnt.rays++; h = bvh_trace<false, false>(sc, ray); si_t = h.t; needs_intersection = false; }
        bool active_medium = medium >= 0, active_surface = !active_medium, escaped_medium = false; float step_t = 0.f;
        V3 tr_multiplier = mk3(1.f);
        if (active_medium) {
            const DMedium &m = sc.media[medium];
            if (medium_uses_dda(sc, m)) {
                // whole medium segment against the local majorants; every tentative collision multiplies by sigma_n / mu
                const DGrid &g = sc.grids[m.m.sigma_t_grid]; float mint, maxt;
                if (medium_clip(g, ray, mint, maxt)) {
                    const float t_end = fminf(maxt, si_t);
                    if (mint < t_end) {
                        DDA dd; dda_init(g, ray, mint, dd); float t = mint, tau = -logf(1.f - smp.next_1d()), mu;
                        while (dda_walk(g, m.m.scale * kPrbMajorantSlack, dd, t, tau, t_end, mu)) {
                            cnt.collisions++;
                            GridTap tap; V3 st = grid_eval<SCATTER>(g, ray_at(ray, t), &tap) * m.m.scale; V3 tm = (mk3(mu) - st) / mu;
                            if (SCATTER) {
                                float gs = 0.f;
                                if (tm.x > 0.f) gs += (dL.x * adj.x / tm.x) * (-m.m.scale / mu);
                                if (tm.y > 0.f) gs += (dL.y * adj.y / tm.y) * (-m.m.scale / mu);
                                if (tm.z > 0.f) gs += (dL.z * adj.z / tm.z) * (-m.m.scale / mu);
                                if (g.differentiable && g.ch == 1 && gs != 0.f) grid_grad_scatter(g, tap, gs, true);
                            }
                            transmittance *= tm;
                            if (!any_nonzero(transmittance)) break;
                            tau = -logf(1.f - smp.next_1d());
                        }
                    }
                }
                escaped_medium = true; active_medium = false;
            } else {
                GridTap tap;
                MediumInteraction mei = sample_interaction<SCATTER>(sc, medium, ray, smp.next_1d(), channel, &tap, cnt.collisions);
                if (si_t < mei.t) mei.t = LR_INF;
                if (sc.prb_nee_homogeneous && m.m.kind == LR_MEDIUM_HOMOGENEOUS) {
                    float t = fminf(remaining_dist, si_t) - mei.mint;
                    tr_multiplier = exp3(mei.combined_extinction * (-t)); mei.t = LR_INF;
                }
                escaped_medium = !mei_valid(mei); active_medium = mei_valid(mei);
                if (active_medium) {
                    ray.o = mei.p; si_t = si_t - mei.t; step_t = mei.t; tr_multiplier *= mei.sigma_n / mei.combined_extinction;
                    if (SCATTER && m.m.kind == LR_MEDIUM_HETEROGENEOUS) {
                        const DGrid &g = sc.grids[m.m.sigma_t_grid]; float gs = 0.f;
                        if (tr_multiplier.x > 0.f) gs += (dL.x * adj.x / tr_multiplier.x) * (-m.m.scale / mei.combined_extinction.x);
                        if (tr_multiplier.y > 0.f) gs += (dL.y * adj.y / tr_multiplier.y) * (-m.m.scale / mei.combined_extinction.y);
                        if (tr_multiplier.z > 0.f) gs += (dL.z * adj.z / tr_multiplier.z) * (-m.m.scale / mei.combined_extinction.z);
                        if (g.differentiable && g.ch == 1 && gs != 0.f) grid_grad_scatter(g, tap, gs, true);
                    }
                }
            }
        }
        active_surface |= escaped_medium;
        active_surface &= (si_t != LR_INF) && !active_medium;
        SurfaceInteraction si;
        if (active_surface) {
            h.t = si_t; si = compute_si(sc, ray, h, a.tris_gid);
            BSDF bsdf(sc, si); tr_multiplier *= bsdf.eval_null_transmission();
        }
        transmittance *= tr_multiplier;
        if (active_surface) { Ray r2 = spawn_ray(si.p, si.n, ray.d); ray.o = r2.o; }
        ray.maxt = remaining_dist;
        needs_intersection |= active_surface;
        active &= (active_medium || active_surface) && any_nonzero(transmittance);
        if (active) total_dist += active_medium ? step_t : si_t;
        if (active_surface && mesh_is_transition(sc, si.mesh)) medium = target_medium(sc, si, ray.d);
    }
    return transmittance;
}

template <bool ADJ>
__global__ void __launch_bounds__(256, LR_WFP_OCC) k_wfp_nee(const __grid_constant__ DScene sc, const __grid_constant__ WFArgs a, uint32_t surf_mask) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0; const float sgn = ADJ ? -1.f : 1.f;
    (void) surf_mask;
    {
        const uint32_t *queue = a.q_nee;
        WF_FOR_QUEUE(a.q->n_nee) {
            if (qi < _n) {
                WState *S = a.st + queue[qi]; uint32_t f = wf_flags(*S);
                if ((f & WF_NEE_ACTIVE) && !(f & WF_DEAD)) {
                    Sampler smp; smp.state = S->rng_state; smp.inc = S->rng_inc; const Sampler smp0 = smp;
                    const uint32_t channel = f & WF_CHANNEL_MASK;
                    V3 T = wfp_nee_walk<false>(sc, a, S, smp, channel, mk3(0.f), mk3(0.f), cnt);
                    float4 w = S->nw_med, e = S->ne_dist, r4 = S->res_depth; float mis = S->ntr_mis.w;
                    V3 contrib = mk3(w.x, w.y, w.z) * mis * (mk3(e.x, e.y, e.z) * T);
                    V3 L = mk3(r4.x, r4.y, r4.z) + contrib * sgn;
                    if (ADJ) {                              // replay with the cloned sampler (prbvolpath.py:283-291)
                        float4 d4 = S->pad0; Sampler s2 = smp0;
                        (void) wfp_nee_walk<true>(sc, a, S, s2, channel, contrib, mk3(d4.x, d4.y, d4.z), cnt);
                    }
                    S->res_depth = make_float4(L.x, L.y, L.z, r4.w);
                    S->rng_state = smp.state; S->rng_inc = smp.inc;
                }
            }
        }
    }
    flush_counters(cnt, 0, a.stats);
}

// SHADE_B, medium (prbvolpath.py:293-309)
__global__ void __launch_bounds__(256, LR_WFP_OCC) k_wfp_shade_b_medium(const __grid_constant__ DScene sc, const __grid_constant__ WFArgs a) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    uint32_t *q_out = a.q_active[a.cur ^ 1], *n_out = &a.q->n_active[a.cur ^ 1];
    WF_FOR_QUEUE(a.q->n_scatter) {
        bool item = qi < _n; uint32_t slot = item ? a.q_scatter[qi] : 0;
        if (item) {
            WState *S = a.st + slot;
            Sampler smp; smp.state = S->rng_state; smp.inc = S->rng_inc;
            uint32_t f = wf_flags(*S) | WF_VALID_RAY;
            const DMedium &m = sc.media[wf_medium(f)]; Phase phase; phase.kind = m.m.phase; phase.g = m.m.phase_g;
            float4 o = S->o_maxt, d = S->d_eta;
            (void) smp.next_1d(); V2 s2 = smp.next_2d();
            float phase_pdf; V3 wo = phase.sample(-mk3(d.x, d.y, d.z), s2, phase_pdf);
            Ray ray; ray.o = mk3(o.x, o.y, o.z); ray.d = mk3(d.x, d.y, d.z); ray.maxt = o.w;
            if (phase_pdf > 0.f) { ray.d = wo; ray.maxt = LR_LARGEST; S->thr_lpdf.w = phase_pdf; }
            cnt.rays++; Hit h = bvh_trace<false, false>(sc, ray);
            S->rng_state = smp.state; S->rng_inc = smp.inc;
            S->o_maxt.w = ray.maxt; S->d_eta = make_float4(ray.d.x, ray.d.y, ray.d.z, d.w); S->hit = wf_hit4(h);
            S->lsp_flags.w = __uint_as_float(f);
        }
        wf_push(q_out, n_out, slot, item);
    }
    flush_counters(cnt, 0, a.stats);
}

// SHADE_B, surface (prbvolpath.py:311-345): a BSDF sample with pdf = 0 ends the path
template <int C, bool TO_FILM, bool ADJ>
__global__ void __launch_bounds__(128, 2 * LR_WFP_OCC) k_wfp_shade_b_surface(const __grid_constant__ DScene sc, const __grid_constant__ WFArgs a, int bucket) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    uint32_t *q_out = a.q_active[a.cur ^ 1], *n_out = &a.q->n_active[a.cur ^ 1];
    WF_FOR_QUEUE(a.q->n_surf[bucket]) {
        bool item = qi < _n; uint32_t slot = item ? a.q_surf[bucket][qi] : 0; bool live = false, ended = false;
        if (item) {
            WState *S = a.st + slot; uint32_t f = wf_flags(*S);
            if (!(f & WF_DEAD)) {
                Sampler smp; smp.state = S->rng_state; smp.inc = S->rng_inc;
                float4 o = S->o_maxt, d = S->d_eta, t = S->thr_lpdf;
                Ray ray; ray.o = mk3(o.x, o.y, o.z); ray.d = mk3(d.x, d.y, d.z); ray.maxt = o.w;
                SurfaceInteraction si = compute_si(sc, ray, wf_hit(S->hit), a.tris_gid);
                BSDF bsdf(sc, si);
                float s1 = smp.next_1d(); V2 s2 = smp.next_2d();
                V3 bsdf_weight; BSDFSample bs = bsdf.sample(si.wi, s1, s2, bsdf_weight);
                S->rng_state = smp.state; S->rng_inc = smp.inc;
                if (!(bs.pdf > 0.f)) {
                    ended = true; float4 r4 = S->res_depth;
                    if (!ADJ) wf_finish<C, TO_FILM>(sc, a, *S, mk3(r4.x, r4.y, r4.z), f);
                } else {
                    live = true;
                    V3 thr = mk3(t.x, t.y, t.z) * bsdf_weight; float eta = d.w * bs.eta, lpdf = t.w;
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
                    S->o_maxt = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.maxt); S->d_eta = make_float4(ray.d.x, ray.d.y, ray.d.z, eta);
                    S->thr_lpdf = make_float4(thr.x, thr.y, thr.z, lpdf); S->hit = wf_hit4(h);
                    S->lsp_flags.w = __uint_as_float(f);
                }
            }
        }
        wf_push(q_out, n_out, slot, live);
        wf_push(a.q_regen, &a.q->n_regen, slot, ended);
    }
    flush_counters(cnt, 0, a.stats);
}

} // namespace lr
