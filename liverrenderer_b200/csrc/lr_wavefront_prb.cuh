// lr_wavefront_prb.cuh -- wavefront `prbvolpath` (reference: src/python/python/ad/integrators/prbvolpath.py:96-444) on the machinery of
// the forward wavefront (lr_wavefront.cuh): path state in the SoA planes, ray queries in the persistent TRACE stage (k_wv_trace is
// shared as it is: closest hit + routing, any-hit / closest shadow rays), russian roulette of the next vertex drawn by the stage
// that spawns the ray, primary rays that leave the scene finished inside GEN, sealed media skip emitter sampling.
//   ADJ = false: primal pass (radiance per lane -> film or per-lane buffer).
//   ADJ = true : adjoint replay -- the slot's radiance starts at the lane's primal result and every contribution is subtracted
//                as it is replayed; each collision scatters
//                    d loss / d grid += w_k * sum_c dL_c * (L_c / weight_c) * d weight_c / d sigma_t * scale
//                into the gradient grid with the warp-aggregated footprint scatter (grid_grad_scatter, lr_prb.cuh).
// Both passes draw the lane's PCG32 stream in the reference's order, so wavefront == per-lane kernel (k_backward) == oracle.
#pragma once
#include "lr_wavefront.cuh"

namespace lr {

// The detached-sampling gradient of a null collision is dL * L * (-scale) / sigma_n: singular where the majorant is tight.
// The PRB passes therefore walk the super-grid with slackened local majorants (sigma_n >= mu / 2); primal `volpath` keeps
// the tight ones.  Both PRB passes use the same factor, so the adjoint replays exactly the primal paths.
static constexpr float kPrbMajorantSlack = 2.f;

struct WFPrbArgs {                     // extra inputs of the PRB passes
    const float *grad_in; const float *weight_img; uint32_t grad_channels;      // adjoint: dLoss/dImage, per-pixel filter-weight sums
    const float *primal_L;            // adjoint: float4 per lane (local lane index - lane_begin), written by the primal pass
};

// prbvolpath.py:141-153: russian roulette at the top of a loop iteration (cap 0.99; the sample is drawn whenever the lane is active)
LR_DEV bool wvp_rr(const DScene &sc, Sampler &smp, V3 &thr, float eta, uint32_t depth) {
    bool active = any_nonzero(thr);
    const float q = fminf(hmax(thr) * sqr(eta), 0.99f);
    const bool perform_rr = depth > (uint32_t) sc.rr_depth;
    if (active) { const float u = smp.next_1d(); active &= (u < q) || !perform_rr; }
    if (perform_rr) thr *= 1.f / q;
    return active;
}

// end of a path: the primal pass writes the lane's radiance; the adjoint has nothing to write
template <int C, bool TO_FILM, bool ADJ>
LR_DEV void wvp_finish(const DScene &sc, const WVArgs &a, uint32_t slot, bool fin, uint32_t f) {
    if (ADJ) return;
    float4 lp = make_float4(0.f, 0.f, 0.f, 0.f); V3 L = mk3(0.f);
    if (fin) { lp = a.p.lane(slot); const float4 r4 = a.p.res(slot); L = mk3(r4.x, r4.y, r4.z); }
    wv_finish<C, TO_FILM>(sc, a, fin, lp, L, f);
}

// ------------------------------------------------------------------------------------------
// GEN (common.py:231-309 sample_rays; adjoint: delta L of the lane = adjoint of splat + develop, common.py:370-400)
// ------------------------------------------------------------------------------------------
template <int C, bool TO_FILM, bool ADJ>
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wvp_gen(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a, const __grid_constant__ WFPrbArgs pa) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0; unsigned long long nsamples = 0;
    const unsigned l = threadIdx.x & 31u; const uint64_t n_total = a.lane_end - a.lane_begin;
    const bool box = sc.sensor.rfilter == LR_RFILTER_BOX, coalesce = a.spp_per_pass >= 4;
    for (;;) {
        unsigned long long base = ~0ull;
        if (l == 0) {
            const uint32_t used = *(volatile uint32_t *) &a.q->n_alloc;
            if (used + a.slack <= a.cap && *(volatile unsigned long long *) a.lane_next < n_total) base = atomicAdd(a.lane_next, 32ull);
        }
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base == ~0ull || base >= n_total) break;
        const bool have = base + l < n_total; const uint64_t li = a.lane_begin + base + l, lane = lane_global(a.map, li);
        LaneSetup L; L.smp.state = L.smp.inc = 0; L.pos = L.sample_pos = mk2(0.f, 0.f); L.ray.o = L.ray.d = mk3(0.f); L.ray.maxt = 0.f;
        bool keep = false, fin = false; Hit h; h.t = LR_INF; h.prim = 0xffffffffu; h.u = h.v = 0.f; uint32_t f = 0; V3 thr = mk3(1.f), Lr = mk3(0.f);
        if (have) {
            seed_sampler(sc, L.smp, a.seed, lane, a.spp_per_pass, 1);
            lane_begin_sample(sc, L, lane, a.spp_per_pass); lane_next_ray(sc, L);
            const uint32_t channel = (uint32_t) fminf(3.f * L.smp.next_1d(), 2.f);           // prbvolpath.py:136-139
            f = channel | WF_HAS_LANE | WF_SPEC_CHAIN;                                        // valid_ray = false, specular_chain = true
            f = wf_set_medium(f, -1);                                                         // :127-128: the sensor is never inside a medium
            nsamples++;
            const bool alive = wvp_rr(sc, L.smp, thr, 1.f, 0u);
            if (alive) { cnt.rays++; h = bvh_trace<false, false>(sc, L.ray); }
            if (!alive) fin = true;
            else if (!hit_valid(h)) {                                                         // left the scene: :233-257 with an invalid si, then the path ends
                fin = true;
                if (sc.env >= 0 && !sc.hide_emitters) Lr = thr * emitter_eval(sc, sc.env, -L.ray.d.z, L.ray.d);
            } else keep = true;
        }
        const V2 fp2 = box ? L.pos : L.sample_pos;
        if (!ADJ) wv_finish<C, TO_FILM>(sc, a, fin, make_float4(__uint_as_float((uint32_t) lane), __uint_as_float((uint32_t) (lane >> 32)), fp2.x, fp2.y), Lr, f);
        uint32_t slot = 0;
        {
            const unsigned m = __ballot_sync(0xffffffffu, keep); uint32_t b0 = 0;
            if (m) {
                if (l == (unsigned) (__ffs(m) - 1)) b0 = atomicAdd(&a.q->n_alloc, (uint32_t) __popc(m));
                b0 = __shfl_sync(0xffffffffu, b0, __ffs(m) - 1);
                slot = b0 + __popc(m & ((1u << l) - 1u));
            }
        }
        if (keep) {
            const WVPlanes &p = a.p; V3 Lin = mk3(0.f);
            if (ADJ) {
                const float4 pl = reinterpret_cast<const float4 *>(pa.primal_L)[li - a.lane_begin]; Lin = mk3(pl.x, pl.y, pl.z);
                V3 dL = mk3(0.f); Footprint fp; const int W = (int) sc.sensor.film_width, H = (int) sc.sensor.film_height;
                if (film_footprint(sc, fp2, coalesce, fp)) {
                    for (int j = 0; j < fp.ny; ++j) { int y = fp.y0 + j; if (y < 0 || y >= H) continue;
                        for (int i = 0; i < fp.nx; ++i) { int x = fp.x0 + i; if (x < 0 || x >= W) continue;
                            float w = __fmul_rn(fp.wy[j], fp.wx[i]); float ws = pa.weight_img[(size_t) y * W + x]; float d = ws == 0.f ? 1.f : ws;
                            const float *g = pa.grad_in + ((size_t) y * W + x) * pa.grad_channels;
                            float s = w / d; dL.x = fmaf(g[0], s, dL.x); dL.y = fmaf(g[1], s, dL.y); dL.z = fmaf(g[2], s, dL.z); } }
                }
                p.adj(slot) = make_float4(dL.x, dL.y, dL.z, 0.f);
            }
            wv_rng_store(p, slot, L.smp);
            p.o(slot) = make_float4(L.ray.o.x, L.ray.o.y, L.ray.o.z, L.ray.maxt);
            p.d(slot) = make_float4(L.ray.d.x, L.ray.d.y, L.ray.d.z, __uint_as_float(f));
            p.thr(slot) = make_float4(thr.x, thr.y, thr.z, __uint_as_float(0u));
            p.hit(slot) = wf_hit4(h);
            p.res(slot) = make_float4(Lin.x, Lin.y, Lin.z, 1.f);
            p.lsp(slot) = make_float4(0.f, 0.f, 0.f, 1.f);
            p.lane(slot) = make_float4(__uint_as_float((uint32_t) lane), __uint_as_float((uint32_t) (lane >> 32)), fp2.x, fp2.y);
        }
        wv_route(sc, a, slot, keep, f, h);
    }
    flush_counters(cnt, nsamples, a.stats);
}

// adjoint of one tentative collision of the main walk (prbvolpath.py:169-204 + the backward of `weight`)
LR_DEV void wvp_scatter_main(const DScene &sc, const DMedium &m, const GridTap &tap, V3 dL, V3 L, V3 weight, V3 dweight) {
    const DGrid &g = sc.grids[m.m.sigma_t_grid];
    float gsum = (dL.x * (L.x / fmaxf(1e-8f, weight.x))) * dweight.x + (dL.y * (L.y / fmaxf(1e-8f, weight.y))) * dweight.y +
                 (dL.z * (L.z / fmaxf(1e-8f, weight.z))) * dweight.z;
    if (g.differentiable && g.ch == 1 && gsum != 0.f) grid_grad_scatter(g, tap, gsum, true);
}

// ------------------------------------------------------------------------------------------
// TRACK (prbvolpath.py:141-204) for paths inside a medium.  The roulette that opens the iteration was drawn by the stage that
// spawned the ray; the reference's further loop iterations (after a null collision) draw theirs here.
// Events: 0 end of path, 1 real scatter, 2 boundary / escape.
// ------------------------------------------------------------------------------------------
template <int C, bool TO_FILM, bool ADJ>
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wvp_track(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    const WVPlanes &p = a.p;
    WF_FOR_QUEUE(a.q->n_track) {
        const bool item = qi < _n; const uint32_t slot = item ? a.q_track[qi] : 0;
        int event = -1; uint32_t f = 0; Hit h0; h0.t = LR_INF; h0.prim = 0xffffffffu; h0.u = h0.v = 0.f;
        if (item) {
            Sampler smp = wv_rng(p, slot);
            float4 t4 = p.o(slot); Ray ray; ray.o = mk3(t4.x, t4.y, t4.z); ray.maxt = t4.w;
            t4 = p.d(slot); ray.d = mk3(t4.x, t4.y, t4.z); f = __float_as_uint(t4.w);
            t4 = p.thr(slot); V3 thr = mk3(t4.x, t4.y, t4.z); uint32_t depth = __float_as_uint(t4.w);
            h0 = wf_hit(p.hit(slot)); float hit_t = h0.t;
            const float4 r4 = p.res(slot); const float eta = r4.w; const V3 Lcur = mk3(r4.x, r4.y, r4.z);
            V3 dL = mk3(0.f); if (ADJ) { const float4 d4 = p.adj(slot); dL = mk3(d4.x, d4.y, d4.z); }
            const uint32_t channel = f & WF_CHANNEL_MASK; const int medium = wf_medium(f);
            const DMedium &m = sc.media[medium]; const bool hetero = m.m.kind == LR_MEDIUM_HETEROGENEOUS;
            for (bool first = true;; first = false) {
                if (!first && !wvp_rr(sc, smp, thr, eta, depth)) { event = 0; break; }
                GridTap tap; V3 weight, dweight = mk3(0.f); bool valid, null_c = false; V3 pt = mk3(0.f); float t_c = 0.f;
                V3 sigma_t, sigma_n, ce;
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
                    cnt.collisions++; pt = ray_at(ray, t_c);
                    sigma_t = grid_eval<ADJ>(g, pt, &tap) * m.m.scale; ce = mk3(mu); sigma_n = ce - sigma_t;
                    weight = mk3(1.f / mu);
                } else {
                    const float u = smp.next_1d();
                    MediumInteraction mei = sample_interaction<ADJ>(sc, medium, ray, u, channel, &tap, cnt.collisions);
                    if (hit_t < mei.t) mei.t = LR_INF;
                    const float t = fminf(mei.t, hit_t) - mei.mint;
                    const V3 tr = exp3(mei.combined_extinction * (-t));
                    const V3 ffp = hit_t < mei.t ? tr : tr * mei.combined_extinction;
                    const float tr_pdf = idx3(ffp, channel);
                    weight = tr_pdf > 0.f ? tr / tr_pdf : mk3(0.f);
                    valid = mei_valid(mei);
                    if (!valid) { thr *= weight; event = 2; break; }
                    pt = mei.p; t_c = mei.t; sigma_t = mei.sigma_t; sigma_n = mei.sigma_n; ce = mei.combined_extinction;
                }
                // a tentative collision at pt: null or real (prbvolpath.py:181-204)
                float scatter_prob = 1.f;
                if (sc.prb_handle_null) { scatter_prob = hmean(sigma_t / ce); null_c = smp.next_1d() >= scatter_prob; }
                if (null_c) {
                    const V3 A = weight; weight *= sigma_n / (1.f - scatter_prob); dweight = A * (-m.m.scale / (1.f - scatter_prob));
                    thr *= weight;
                    if (ADJ && hetero) wvp_scatter_main(sc, m, tap, dL, Lcur, weight, dweight);
                    ray.o = pt; hit_t = hit_t - t_c;
                    if (!(depth < sc.max_depth)) { event = 0; break; }      // :189-190: `active &= depth < max_depth` ends the lane after this collision
                    continue;
                }
                depth += 1; cnt.real_scatters++;
                { float *lp = reinterpret_cast<float *>(&p.lsp(slot)); lp[0] = pt.x; lp[1] = pt.y; lp[2] = pt.z; }
                if (!(depth < sc.max_depth)) { thr *= weight; event = 0; break; }
                {
                    const V3 A = weight; const V3 albedo = medium_albedo(sc, m, pt);
                    weight *= (sigma_t * albedo) / scatter_prob; dweight = A * albedo * (m.m.scale / scatter_prob);
                    thr *= weight;
                    if (ADJ && hetero) wvp_scatter_main(sc, m, tap, dL, Lcur, weight, dweight);
                    if (ADJ) albedo_grad_scatter(sc, m, pt, albedo_grad_value(dL, Lcur, weight, A * sigma_t / scatter_prob));
                }
                ray.o = pt; hit_t = hit_t - t_c; event = 1; break;
            }
            wv_rng_store(p, slot, smp);
            if (event > 0) {
                p.o(slot) = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.maxt);
                p.thr(slot) = make_float4(thr.x, thr.y, thr.z, __uint_as_float(depth));
                p.hit(slot).x = hit_t;
            }
        }
        wvp_finish<C, TO_FILM, ADJ>(sc, a, slot, event == 0, f);
        const bool esc = event == 2 && !hit_valid(h0);
        const uint32_t b = (event == 2 && !esc) ? wv_bucket(sc, h0) : 0xffu;
        wf_push(a.q_scatter, &a.q->n_scatter, slot, event == 1);
        wf_push(a.q_escape, &a.q->n_escape, slot, esc);
#pragma unroll
        for (uint32_t k = 0; k < WF_N_SURF; ++k) wf_push(a.q_surf[k], &a.q->n_surf[k], slot, b == k);
    }
    flush_counters(cnt, 0, a.stats);
}

// TRACK for scenes whose media all use the majorant DDA (production mode): the per-item work of k_wvp_track as a persistent
// "while-while" loop with the refill scheme of k_wv_track_dda -- a lane whose item has finished parks, and once half of the warp is
// parked the parked lanes are written out and pull the next queue entries.  After a null collision the walk CONTINUES in its DDA state
// (k_wvp_track restarts it from the collision point: the same cell, the same majorant); draw order per lane is unchanged: optical
// depth, then per tentative collision (null test, roulette of the next loop iteration, next optical depth).
// scatter vertex inside a sealed medium: k_wvp_shade_a_medium + k_wvp_shade_b_medium done by TRACK itself (see wv_track_scatter_sealed,
// lr_wavefront.cuh): flags, the two discarded emitter draws, phase sample, roulette of the next iteration -> TRACE_C (event 3) or the end (0)
__noinline__ LR_DEV void wvp_track_scatter_sealed(const DScene &sc, const WVPlanes &p, uint32_t slot, Sampler &smp, Ray &ray, V3 &thr, float eta, uint32_t depth, uint32_t &f, int &event) {
    const DMedium &m = sc.media[wf_medium(f)];
    if (!m.sealed) return;
    const bool sample_emitters = m.m.sample_emitters != 0;
    f = sample_emitters ? (f & ~WF_SPEC_CHAIN) : (f | WF_SPEC_CHAIN); f &= ~(WF_NEE_ACTIVE | WF_DEAD);
    if (sample_emitters) (void) smp.next_2d();
    f |= WF_VALID_RAY;
    Phase phase; phase.kind = m.m.phase; phase.g = m.m.phase_g;
    (void) smp.next_1d(); const V2 s2 = smp.next_2d();
    float phase_pdf; const V3 wo = phase.sample(-ray.d, s2, phase_pdf);
    if (wvp_rr(sc, smp, thr, depth > (uint32_t) sc.rr_depth ? eta : 1.f, depth)) {
        if (phase_pdf > 0.f) { ray.d = wo; ray.maxt = LR_LARGEST; reinterpret_cast<float *>(&p.lsp(slot))[3] = phase_pdf; }
        event = 3;
    } else event = 0;
}

template <int C, bool TO_FILM, bool ADJ>
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wvp_track_dda(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a, const int fuse) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    const WVPlanes &p = a.p;
    const uint32_t n_items = a.q->n_track;
    const unsigned lane = threadIdx.x & 31u, lt = (1u << lane) - 1u;
    uint32_t slot = 0, depth = 0, f = 0; int event = -2;                  // -2 idle, -1 walking, 0 end of path, 1 scatter, 2 boundary
    Sampler smp; smp.state = smp.inc = 0; Ray ray; ray.o = ray.d = mk3(0.f); ray.maxt = 0.f; V3 thr = mk3(0.f), Lcur = mk3(0.f), dL = mk3(0.f);
    float hit_t = 0.f, t = 0.f, tau = 0.f, t_end = 0.f, eta = 1.f; int grid_id = 0, medium = 0; DDA dd{};
    Hit h0; h0.t = LR_INF; h0.prim = 0xffffffffu; h0.u = h0.v = 0.f;
    bool exhausted = false;
    for (;;) {
        const unsigned walking = __ballot_sync(0xffffffffu, event == -1);
        if (__popc(walking) <= LR_TRACK_REFILL) {
            // ---- write finished items out and queue them (converged: the pushes use warp ballots) ----
            if (__any_sync(0xffffffffu, event >= 0)) {
                if (event == 1 && fuse) wvp_track_scatter_sealed(sc, p, slot, smp, ray, thr, eta, depth, f, event);
                if (event >= 0) wv_rng_store(p, slot, smp);
                if (event > 0) {
                    p.o(slot) = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.maxt);
                    p.thr(slot) = make_float4(thr.x, thr.y, thr.z, __uint_as_float(depth));
                    if (event == 3) p.d(slot) = make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float(f));
                    else p.hit(slot).x = hit_t;
                }
                wvp_finish<C, TO_FILM, ADJ>(sc, a, slot, event == 0, f);
                wf_push(a.q_trace, &a.q->n_trace, slot, event == 3);
                const bool esc = event == 2 && !hit_valid(h0);
                const uint32_t b = (event == 2 && !esc) ? wv_bucket(sc, h0) : 0xffu;
                wf_push(a.q_scatter, &a.q->n_scatter, slot, event == 1);
                wf_push(a.q_escape, &a.q->n_escape, slot, esc);
#pragma unroll
                for (uint32_t k = 0; k < WF_N_SURF; ++k) wf_push(a.q_surf[k], &a.q->n_surf[k], slot, b == k);
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
                    t4 = p.res(slot); eta = t4.w; Lcur = mk3(t4.x, t4.y, t4.z);
                    if (ADJ) { t4 = p.adj(slot); dL = mk3(t4.x, t4.y, t4.z); }
                    medium = wf_medium(f); const DMedium &m = sc.media[medium]; grid_id = m.m.sigma_t_grid; const DGrid &g = sc.grids[grid_id];
                    event = 2;
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
            const DGrid &g = sc.grids[grid_id]; const float mu_scale = sc.media[medium].m.scale * kPrbMajorantSlack;
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
        // ---- tentative collisions, evaluated together (prbvolpath.py:181-204) ----
        if (collide) {
            const DGrid &g = sc.grids[grid_id]; const DMedium &m = sc.media[medium];
            cnt.collisions++;
            const V3 pt = ray_at(ray, t);
            GridTap tap; const V3 sigma_t = grid_eval<ADJ>(g, pt, &tap) * m.m.scale, ce = mk3(mu), sigma_n = ce - sigma_t;
            V3 weight = mk3(1.f / mu), dweight;
            float scatter_prob = 1.f; bool null_c = false;
            if (sc.prb_handle_null) { scatter_prob = hmean(sigma_t / ce); null_c = smp.next_1d() >= scatter_prob; }
            if (null_c) {
                const V3 A = weight; weight *= sigma_n / (1.f - scatter_prob); dweight = A * (-m.m.scale / (1.f - scatter_prob));
                thr *= weight;
                if (ADJ) wvp_scatter_main(sc, m, tap, dL, Lcur, weight, dweight);
                if (!(depth < sc.max_depth)) event = 0;                      // :189-190
                else if (!wvp_rr(sc, smp, thr, eta, depth)) event = 0;       // roulette of the next loop iteration
                else tau = -logf(1.f - smp.next_1d());
            } else {
                depth += 1; cnt.real_scatters++;
                { float *lp = reinterpret_cast<float *>(&p.lsp(slot)); lp[0] = pt.x; lp[1] = pt.y; lp[2] = pt.z; }
                if (!(depth < sc.max_depth)) { thr *= weight; event = 0; }
                else {
                    const V3 A = weight; const V3 albedo = medium_albedo(sc, m, pt);
                    weight *= (sigma_t * albedo) / scatter_prob; dweight = A * albedo * (m.m.scale / scatter_prob);
                    thr *= weight;
                    if (ADJ) { wvp_scatter_main(sc, m, tap, dL, Lcur, weight, dweight); albedo_grad_scatter(sc, m, pt, albedo_grad_value(dL, Lcur, weight, A * sigma_t / scatter_prob)); }
                    ray.o = pt; hit_t = hit_t - t; event = 1;
                }
            }
        }
        if (exhausted && !__any_sync(0xffffffffu, event != -2)) break;
    }
    flush_counters(cnt, 0, a.stats);
}

// ------------------------------------------------------------------------------------------
// SHADE_A: emitter sampling (prbvolpath.py:354-378).  The shadow ray goes to the NEE planes; TRACE_S / NEE finish the estimate.
// `nw` holds sgn * throughput * (phase | bsdf) value * MIS weight * emitter value: +1 primal, -1 adjoint (contributions are subtracted).
// ------------------------------------------------------------------------------------------
LR_DEV bool wvp_nee_setup(const DScene &sc, const WVArgs &a, uint32_t slot, Sampler &smp, V3 ref_p, V3 ref_n, const SurfaceInteraction *ref_si,
                          int &medium, DirectionSample &ds, V3 &emitter_val) {
    // production tracking: a shadow ray that starts in a sealed medium always ends on an opaque surface (DMedium::sealed) => exactly 0,
    // value and gradient; the two draws of the emitter sample are consumed as usual
    const bool dda = sc.integ.tracking == LR_TRACKING_DDA;
    if (dda && !ref_si && medium >= 0 && sc.media[medium].sealed) { (void) smp.next_2d(); ds.pdf = 0.f; return false; }
    ds = sample_emitter_direction(sc, ref_p, smp.next_2d(), emitter_val);
    if (ds.pdf == 0.f) return false;
    if (ref_si && mesh_is_transition(sc, ref_si->mesh)) medium = target_medium(sc, *ref_si, ds.d);
    if (dda && medium >= 0 && sc.media[medium].sealed) return false;
    const Ray ray = spawn_ray_to(ref_p, ref_n, ds.p);
    a.p.no(slot) = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.maxt);
    a.p.nd(slot) = make_float4(ray.d.x, ray.d.y, ray.d.z, ds.dist);
    return true;
}

// medium scatter (prbvolpath.py:259-291, medium part)
template <bool ADJ>
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wvp_shade_a_medium(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a) {
    const WVPlanes &p = a.p; const float sgn = ADJ ? -1.f : 1.f;
    WF_FOR_QUEUE(a.q->n_scatter) {
        const bool item = qi < _n; bool nee = false; const uint32_t slot = item ? a.q_scatter[qi] : 0;
        if (item) {
            const float4 d4 = p.d(slot); uint32_t f = __float_as_uint(d4.w); const uint32_t f0 = f; const int medium = wf_medium(f);
            const DMedium &m = sc.media[medium]; Phase phase; phase.kind = m.m.phase; phase.g = m.m.phase_g;
            const bool sample_emitters = m.m.sample_emitters != 0;
            f = sample_emitters ? (f & ~WF_SPEC_CHAIN) : (f | WF_SPEC_CHAIN); f &= ~(WF_NEE_ACTIVE | WF_DEAD);
            if (sample_emitters) {
                Sampler smp = wv_rng(p, slot);
                const float4 o4 = p.o(slot), t4 = p.thr(slot); const V3 pt = mk3(o4.x, o4.y, o4.z), wi = -mk3(d4.x, d4.y, d4.z), thr = mk3(t4.x, t4.y, t4.z);
                DirectionSample ds; V3 ev; int walk_medium = medium;
                if (wvp_nee_setup(sc, a, slot, smp, pt, mk3(0.f), nullptr, walk_medium, ds, ev)) {
                    const float pv = phase.eval_pdf(wi, ds.d);
                    const V3 w = (thr * pv) * mis_weight(ds.pdf, ds.delta ? 0.f : pv) * ev * sgn;
                    p.nw(slot) = make_float4(w.x, w.y, w.z, __int_as_float(walk_medium));
                    nee = true;
                }
                wv_rng_store(p, slot, smp);
            }
            if (f != f0) reinterpret_cast<float *>(&p.d(slot))[3] = __uint_as_float(f);
        }
        wf_push(a.q_shadow, &a.q->n_shadow, slot, nee);
    }
}

// surface (prbvolpath.py:206-291, surface part); every item has a valid hit (escaped rays have their own queue)
template <int C, bool TO_FILM, bool ADJ>
__global__ void __launch_bounds__(128, LR_WF_OCC_SURF) k_wvp_shade_a_surface(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a, int bucket) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0; const float sgn = ADJ ? -1.f : 1.f;
    const WVPlanes &p = a.p;
    WF_FOR_QUEUE(a.q->n_surf[bucket]) {
        const bool item = qi < _n; const uint32_t slot = item ? a.q_surf[bucket][qi] : 0; bool nee = false, gone = false; uint32_t f = 0;
        if (item) {
            Sampler smp = wv_rng(p, slot);
            const float4 o4 = p.o(slot), d4 = p.d(slot), t4 = p.thr(slot);
            f = __float_as_uint(d4.w) & ~(WF_NEE_ACTIVE | WF_DEAD); const uint32_t f0 = __float_as_uint(d4.w);
            const uint32_t depth = __float_as_uint(t4.w); const int medium = wf_medium(f);
            Ray ray; ray.o = mk3(o4.x, o4.y, o4.z); ray.d = mk3(d4.x, d4.y, d4.z); ray.maxt = o4.w;
            const V3 thr = mk3(t4.x, t4.y, t4.z);
            Hit h = wf_hit(p.hit(slot));
            SurfaceInteraction si = compute_si(sc, ray, h, sc.tris_gid);
            if (sc.hide_emitters && medium < 0 && depth == 0u && sc.meshes[si.mesh].emitter >= 0) {                        // :218-231
                Ray r2 = spawn_ray(si.p, si.n, ray.d); Hit pi = skip_area_emitters(sc, sc.tris_gid, r2, cnt);
                si = compute_si(sc, ray, pi, sc.tris_gid); h = pi; p.hit(slot) = wf_hit4(h);
            }
            const bool count_direct = depth == 0u || (f & WF_SPEC_CHAIN);
            const int emitter = emitter_of(sc, si);
            if (emitter >= 0 && !(depth == 0u && sc.hide_emitters)) {                                                      // :233-257
                const float4 l4 = p.lsp(slot); DirectionSample ds = ds_from_si(sc, si, mk3(l4.x, l4.y, l4.z));
                const float emitter_pdf = pdf_emitter_direction(sc, ds);
                const V3 emitted = eval_emitter_at(sc, emitter, si);
                const V3 contrib = count_direct ? thr * emitted : thr * mis_weight(l4.w, emitter_pdf) * emitted;
                float4 r4 = p.res(slot); r4.x += contrib.x * sgn; r4.y += contrib.y * sgn; r4.z += contrib.z * sgn; p.res(slot) = r4;
            }
            // a lane that arrives with depth == max_depth still counts the emitter it sees (`active_surface` is not re-masked
            // after `active &= depth < max_depth`, prbvolpath.py:189-257) and ends here
            if (!si_valid(si) || !(depth < sc.max_depth)) { gone = true; f |= WF_DEAD; }
            else {
                BSDF bsdf(sc, si);
                if ((bsdf.flags() & F_SMOOTH) && (depth + 1 < sc.max_depth)) {
                    DirectionSample ds; V3 ev; int walk_medium = medium;
                    if (wvp_nee_setup(sc, a, slot, smp, si.p, si.n, &si, walk_medium, ds, ev)) {
                        float nee_pdf; const V3 bv = bsdf.eval_pdf(si.wi, si.sh.to_local(ds.d), nee_pdf);
                        const V3 w = (thr * bv) * mis_weight(ds.pdf, ds.delta ? 0.f : nee_pdf) * ev * sgn;
                        p.nw(slot) = make_float4(w.x, w.y, w.z, __int_as_float(walk_medium));
                        nee = true;
                    }
                    wv_rng_store(p, slot, smp);
                }
            }
            if (f != f0) reinterpret_cast<float *>(&p.d(slot))[3] = __uint_as_float(f);
        }
        wvp_finish<C, TO_FILM, ADJ>(sc, a, slot, gone, f);
        wf_push(a.q_shadow, &a.q->n_shadow, slot, nee);
    }
    flush_counters(cnt, 0, a.stats);
}

// rays that left the scene (prbvolpath.py:233-257 with an invalid si): environment radiance + MIS, then the path ends
template <int C, bool TO_FILM, bool ADJ>
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wvp_escape(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a) {
    const WVPlanes &p = a.p;
    if (ADJ) return;                                         // the adjoint of an emitter hit carries no grid gradient: nothing to replay
    WF_FOR_QUEUE(a.q->n_escape) {
        const bool item = qi < _n; const uint32_t slot = item ? a.q_escape[qi] : 0; uint32_t f = 0;
        if (item) {
            const float4 d4 = p.d(slot), t4 = p.thr(slot);
            f = __float_as_uint(d4.w); const uint32_t depth = __float_as_uint(t4.w);
            if (sc.env >= 0 && !(depth == 0u && sc.hide_emitters)) {
                const V3 dir = mk3(d4.x, d4.y, d4.z), thr = mk3(t4.x, t4.y, t4.z);
                const bool count_direct = depth == 0u || (f & WF_SPEC_CHAIN);
                const V3 emitted = emitter_eval(sc, sc.env, -dir.z, dir);
                V3 contrib;
                if (count_direct) contrib = thr * emitted;
                else {
                    DirectionSample ds; ds.d = dir; ds.emitter = sc.env; ds.dist = 0.f; ds.n = ds.p = mk3(0.f); ds.uv = mk2(0.f, 0.f); ds.pdf = 0.f; ds.delta = false;
                    contrib = thr * mis_weight(p.lsp(slot).w, pdf_emitter_direction(sc, ds)) * emitted;
                }
                float4 r4 = p.res(slot); r4.x += contrib.x; r4.y += contrib.y; r4.z += contrib.z; p.res(slot) = r4;
            }
        }
        wvp_finish<C, TO_FILM, false>(sc, a, slot, item, f);
    }
}

// ------------------------------------------------------------------------------------------
// NEE: one shadow-segment walk of PRBVolpathIntegrator.sample_emitter (prbvolpath.py:379-444) over the slots TRACE_S let through.
// SCATTER = false: returns the transmittance.  SCATTER = true: replays the walk (same sampler state) and back-propagates `adj` =
// the NEE contribution through every tr_multiplier = sigma_n / majorant:  d/d grid = dL * adj / tr_mult * (-scale / majorant).
// ------------------------------------------------------------------------------------------
template <bool SCATTER>
LR_DEV V3 wvp_nee_walk(const DScene &sc, const WVArgs &a, uint32_t slot, bool have_nhit, Sampler &smp, uint32_t channel, V3 adj, V3 dL, Counters &cnt) {
    const WVPlanes &p = a.p;
    const float4 o = p.no(slot), d = p.nd(slot); Ray ray; ray.o = mk3(o.x, o.y, o.z); ray.d = mk3(d.x, d.y, d.z);
    const float max_dist = o.w; float total_dist = 0.f;
    int medium = __float_as_int(p.nw(slot).w);
    Hit h; h.t = LR_INF; h.prim = 0xffffffffu; h.u = h.v = 0.f;
    if (have_nhit) h = wf_hit(p.nhit(slot));
    float si_t = h.t; bool needs_intersection = false, active = true;
    V3 transmittance = mk3(1.f);
    while (active) {
        const float remaining_dist = max_dist - total_dist; ray.maxt = remaining_dist;
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
                            GridTap tap; const V3 st = grid_eval<SCATTER>(g, ray_at(ray, t), &tap) * m.m.scale; const V3 tm = (mk3(mu) - st) / mu;
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
                    const float t = fminf(remaining_dist, si_t) - mei.mint;
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
            h.t = si_t; si = compute_si(sc, ray, h, sc.tris_gid);
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
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wvp_nee(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a, int have_nhit) {
    Counters cnt; cnt.rays = cnt.collisions = cnt.real_scatters = 0;
    const WVPlanes &p = a.p;
    WF_FOR_QUEUE(a.q->n_nee) {
        if (qi < _n) {
            const uint32_t slot = a.q_nee[qi];
            Sampler smp = wv_rng(p, slot); const Sampler smp0 = smp;
            const uint32_t channel = __float_as_uint(p.d(slot).w) & WF_CHANNEL_MASK;
            const V3 T = wvp_nee_walk<false>(sc, a, slot, have_nhit != 0, smp, channel, mk3(0.f), mk3(0.f), cnt);
            const float4 w4 = p.nw(slot); const V3 cs = mk3(w4.x, w4.y, w4.z) * T;          // signed contribution
            float4 r4 = p.res(slot); r4.x += cs.x; r4.y += cs.y; r4.z += cs.z; p.res(slot) = r4;
            if (ADJ) {                                  // replay with the cloned sampler (prbvolpath.py:283-291); adj = the (unsigned) contribution
                const float4 d4 = p.adj(slot); Sampler s2 = smp0;
                (void) wvp_nee_walk<true>(sc, a, slot, have_nhit != 0, s2, channel, -cs, mk3(d4.x, d4.y, d4.z), cnt);
            }
            wv_rng_store(p, slot, smp);
        }
    }
    flush_counters(cnt, 0, a.stats);
}

// ------------------------------------------------------------------------------------------
// SHADE_B: direction sampling, then the russian roulette of the next loop iteration.
// ------------------------------------------------------------------------------------------
// medium scatter: phase-function sampling (prbvolpath.py:293-309)
template <int C, bool TO_FILM, bool ADJ>
__global__ void __launch_bounds__(256, LR_WF_OCC) k_wvp_shade_b_medium(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a) {
    const WVPlanes &p = a.p;
    WF_FOR_QUEUE(a.q->n_scatter) {
        const bool item = qi < _n; const uint32_t slot = item ? a.q_scatter[qi] : 0; bool alive = false; uint32_t f = 0;
        if (item) {
            Sampler smp = wv_rng(p, slot);
            const float4 d4 = p.d(slot), t4 = p.thr(slot); f = __float_as_uint(d4.w) | WF_VALID_RAY;
            const DMedium &m = sc.media[wf_medium(f)]; Phase phase; phase.kind = m.m.phase; phase.g = m.m.phase_g;
            (void) smp.next_1d(); const V2 s2 = smp.next_2d();
            float phase_pdf; const V3 wo = phase.sample(-mk3(d4.x, d4.y, d4.z), s2, phase_pdf);
            V3 thr = mk3(t4.x, t4.y, t4.z); const uint32_t depth = __float_as_uint(t4.w);
            const float eta = depth > (uint32_t) sc.rr_depth ? p.res(slot).w : 1.f;
            alive = wvp_rr(sc, smp, thr, eta, depth);
            wv_rng_store(p, slot, smp);
            if (alive) {
                if (phase_pdf > 0.f) {
                    p.d(slot) = make_float4(wo.x, wo.y, wo.z, __uint_as_float(f));
                    reinterpret_cast<float *>(&p.o(slot))[3] = LR_LARGEST; reinterpret_cast<float *>(&p.lsp(slot))[3] = phase_pdf;
                } else reinterpret_cast<float *>(&p.d(slot))[3] = __uint_as_float(f);
                p.thr(slot) = make_float4(thr.x, thr.y, thr.z, t4.w);
            }
        }
        wvp_finish<C, TO_FILM, ADJ>(sc, a, slot, item && !alive, f);
        wf_push(a.q_trace, &a.q->n_trace, slot, alive);
    }
}

// surface: BSDF sampling, next ray, medium transition (prbvolpath.py:311-345): a BSDF sample with pdf = 0 ends the path
template <int C, bool TO_FILM, bool ADJ>
__global__ void __launch_bounds__(128, LR_WF_OCC_SURF) k_wvp_shade_b_surface(const __grid_constant__ DScene sc, const __grid_constant__ WVArgs a, int bucket) {
    const WVPlanes &p = a.p;
    WF_FOR_QUEUE(a.q->n_surf[bucket]) {
        bool item = qi < _n; const uint32_t slot = item ? a.q_surf[bucket][qi] : 0; bool alive = false; uint32_t f = 0;
        float4 d4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (item) { d4 = p.d(slot); f = __float_as_uint(d4.w); item = !(f & WF_DEAD); }        // WF_DEAD: finished by SHADE_A
        if (item) {
            Sampler smp = wv_rng(p, slot);
            const float4 o4 = p.o(slot), t4 = p.thr(slot), r4 = p.res(slot);
            Ray ray; ray.o = mk3(o4.x, o4.y, o4.z); ray.d = mk3(d4.x, d4.y, d4.z); ray.maxt = o4.w;
            const SurfaceInteraction si = compute_si(sc, ray, wf_hit(p.hit(slot)), sc.tris_gid);
            BSDF bsdf(sc, si);
            const float s1 = smp.next_1d(); const V2 s2 = smp.next_2d();
            V3 bsdf_weight; const BSDFSample bs = bsdf.sample(si.wi, s1, s2, bsdf_weight);
            if (bs.pdf > 0.f) {
                V3 thr = mk3(t4.x, t4.y, t4.z) * bsdf_weight; const float eta = r4.w * bs.eta; uint32_t depth = __float_as_uint(t4.w);
                ray = spawn_ray(si.p, si.n, si.sh.to_world(bs.wo));
                const bool non_null = !(bs.sampled_type & F_NULL);
                if (non_null) { depth += 1u; p.lsp(slot) = make_float4(si.p.x, si.p.y, si.p.z, bs.pdf); f |= WF_VALID_RAY; }
                bool spec = (f & WF_SPEC_CHAIN) != 0;
                spec |= non_null && (bs.sampled_type & F_DELTA); spec &= !(bs.sampled_type & F_SMOOTH);
                f = spec ? (f | WF_SPEC_CHAIN) : (f & ~WF_SPEC_CHAIN);
                if (mesh_is_transition(sc, si.mesh)) f = wf_set_medium(f, target_medium(sc, si, ray.d));
                alive = wvp_rr(sc, smp, thr, eta, depth);
                if (alive) {
                    p.o(slot) = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.maxt);
                    p.d(slot) = make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float(f));
                    p.thr(slot) = make_float4(thr.x, thr.y, thr.z, __uint_as_float(depth));
                    if (eta != r4.w) reinterpret_cast<float *>(&p.res(slot))[3] = eta;
                }
            }
            wv_rng_store(p, slot, smp);
        }
        wvp_finish<C, TO_FILM, ADJ>(sc, a, slot, item && !alive, f);
        wf_push(a.q_trace, &a.q->n_trace, slot, alive);
    }
}

} // namespace lr
