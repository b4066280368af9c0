// lr_prb.cuh -- path replay backpropagation for volpath (device code).
// PRBVolpathIntegrator.sample / .sample_emitter (src/python/python/ad/integrators/prbvolpath.py:96-444).
// The adjoint pass replays the primal path with the same PCG32 stream and scatters
//   d(loss)/d(grid voxel k) += w_k * sum_c dL_c * d(weight_c)/d(sigma_t grid value) * Lo_c
// per collision: the closed form of what the reference obtains by dr.backward() through
// `weight` (:169-204) and `tr_multiplier` (:395-427).  sigma_t = scale*grid(p),
// sigma_n = majorant - sigma_t with the majorant detached (src/media/heterogeneous.cpp:163).
#pragma once
#include "lr_integrators.cuh"

namespace lr {

// Warp-aggregated scatter-add of one trilinear footprint into the gradient grid.
// Lanes of a warp that hit the same base voxel (match.any on the 000-corner index with identical
// footprint topology) first reduce their contributions with shuffles; one leader issues the
// 8 RED.ADD.F32.  In dense regions neighbouring paths collide in the same cell all the time.
LR_DEV void grid_grad_scatter(const DGrid &g, const GridTap &tap, float gval, bool lane_active) {
    unsigned active_mask = __ballot_sync(__activemask(), lane_active);
    if (!lane_active) return;
    // key: corner 000 and corner 111 indices identify the whole footprint
    unsigned long long key = ((unsigned long long) tap.idx[0] << 32) | tap.idx[7];
    unsigned peers = __match_any_sync(active_mask, key);
    int leader = __ffs(peers) - 1; int me = threadIdx.x & 31;
    float c[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) c[k] = gval * tap.w[k];
    if (peers != (1u << me)) {
        // segmented butterfly over the peer set (ranks within `peers`): in round `step` the peer of rank r, r a multiple of 2 * step,
        // adds the partial sum of rank r + step; after ceil(log2 n) rounds rank 0 = the leader holds the total.  (The loop this
        // replaces had every peer walk the whole set: n - 1 rounds of 8 shuffles.)
        const int n = __popc(peers), rank = __popc(peers & ((1u << me) - 1u));
        for (int step = 1; step < n; step <<= 1) {
            const unsigned src = __fns(peers, (unsigned) me, step + 1);                 // lane of rank + step, 0xffffffff beyond the set
            const bool take = (rank & (2 * step - 1)) == 0 && src != 0xffffffffu;
            const int from = src != 0xffffffffu ? (int) src : me;
#pragma unroll
            for (int k = 0; k < 8; ++k) { const float v = __shfl_sync(peers, c[k], from); if (take) c[k] += v; }
        }
    }
    if (me == leader) {
#pragma unroll
        for (int k = 0; k < 8; ++k) if (c[k] != 0.f) atomicAdd(g.grad + tap.idx[k], c[k]);
    }
}

// Gradient w.r.t. an ALBEDO grid: sigma_s = sigma_t * albedo (heterogeneous.cpp:189-195), so at a real scatter
// d weight_c / d albedo_c = A_c * sigma_t_c / scatter_prob (prbvolpath.py:202-204 back-propagated through mei.sigma_s).
// `gv`: per channel dL * (L / weight) * d weight / d albedo.  1- or 3-channel grid; real scatters are ~1/10 of the tentative collisions, so
// the 8 x ch atomics are issued directly.
LR_DEV void albedo_grad_scatter(const DScene &sc, const DMedium &m, V3 p, V3 gv) {
    if (m.m.albedo_grid < 0) return;
    const DGrid &g = sc.grids[m.m.albedo_grid];
    if (!g.differentiable || (gv.x == 0.f && gv.y == 0.f && gv.z == 0.f)) return;
    GridTap tap; (void) grid_eval<true>(g, p, &tap);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        if (tap.w[k] == 0.f) continue;
        if (g.ch == 1) atomicAdd(g.grad + tap.idx[k], (gv.x + gv.y + gv.z) * tap.w[k]);
        else { float *d = g.grad + (size_t) tap.idx[k] * 3; atomicAdd(d, gv.x * tap.w[k]); atomicAdd(d + 1, gv.y * tap.w[k]); atomicAdd(d + 2, gv.z * tap.w[k]); }
    }
}
LR_DEV V3 albedo_grad_value(V3 dL, V3 L, V3 weight, V3 dweight_albedo) {
    return mk3((dL.x * (L.x / fmaxf(1e-8f, weight.x))) * dweight_albedo.x, (dL.y * (L.y / fmaxf(1e-8f, weight.y))) * dweight_albedo.y,
               (dL.z * (L.z / fmaxf(1e-8f, weight.z))) * dweight_albedo.z);
}

// PRBVolpathIntegrator.sample_emitter (:354-444).  ADJ: replay with adj_emitted / dL.
template <bool ADJ>
LR_DEV V3 sample_emitter_prb(const DScene &sc, const DTri *tris_gid, V3 ref_p, V3 ref_n, const SurfaceInteraction *ref_si, Sampler &smp,
                             int medium, uint32_t channel, DirectionSample &ds, V3 adj, V3 dL, Counters &cnt) {
    V3 emitter_val;
    ds = sample_emitter_direction(sc, ref_p, smp.next_2d(), emitter_val);
    bool active = ds.pdf != 0.f;
    if (!active) medium = -1;
    if (ref_si && mesh_is_transition(sc, ref_si->mesh)) medium = target_medium(sc, *ref_si, ds.d);
    if (!active) return mk3(0.f);
    Ray ray = spawn_ray_to(ref_p, ref_n, ds.p);
    float max_dist = ray.maxt, total_dist = 0.f; SurfaceInteraction si = invalid_si(); bool needs_intersection = true; V3 transmittance = mk3(1.f);
    while (active) {
        float remaining_dist = max_dist - total_dist; ray.maxt = remaining_dist;
        active &= remaining_dist > 0.f;
        if (!active) break;
        if (needs_intersection) si = intersect(sc, tris_gid, ray, cnt);
        needs_intersection = false;
        bool active_medium = medium >= 0, active_surface = !active_medium, escaped_medium = false;
        MediumInteraction mei; mei.t = LR_INF; V3 tr_multiplier = mk3(1.f); GridTap tap;
        if (active_medium) {
            const DMedium &m = sc.media[medium];
            mei = sample_interaction<ADJ>(sc, medium, ray, smp.next_1d(), channel, &tap, cnt.collisions);
            if (si.t < mei.t) mei.t = LR_INF;
            if (sc.prb_nee_homogeneous && m.m.kind == LR_MEDIUM_HOMOGENEOUS) {
                float t = fminf(remaining_dist, si.t) - mei.mint;
                tr_multiplier = exp3(mei.combined_extinction * (-t)); mei.t = LR_INF;
            }
            escaped_medium = !mei_valid(mei); active_medium = mei_valid(mei);
            if (active_medium) { ray.o = mei.p; si.t = si.t - mei.t; tr_multiplier *= mei.sigma_n / mei.combined_extinction; }
        }
        active_surface |= escaped_medium;
        active_surface &= si_valid(si) && !active_medium;
        if (active_surface) { BSDF bsdf(sc, si); tr_multiplier *= bsdf.eval_null_transmission(); }
        if (ADJ) {
            bool do_adj = active_medium && sc.media[medium].m.kind == LR_MEDIUM_HETEROGENEOUS;
            float gsum = 0.f;
            if (do_adj) {
                const DMedium &m = sc.media[medium];
                // d tr_multiplier_c / d grid = -scale / majorant_c ; backward(tr_mult * dL*adj/tr_mult)
                if (tr_multiplier.x > 0.f) gsum += (dL.x * adj.x / tr_multiplier.x) * (-m.m.scale / mei.combined_extinction.x);
                if (tr_multiplier.y > 0.f) gsum += (dL.y * adj.y / tr_multiplier.y) * (-m.m.scale / mei.combined_extinction.y);
                if (tr_multiplier.z > 0.f) gsum += (dL.z * adj.z / tr_multiplier.z) * (-m.m.scale / mei.combined_extinction.z);
                const DGrid &g = sc.grids[m.m.sigma_t_grid];
                do_adj = g.differentiable && g.ch == 1 && gsum != 0.f;
                if (do_adj) grid_grad_scatter(g, tap, gsum, true);
            }
        }
        transmittance *= tr_multiplier;
        if (active_surface) { Ray r2 = spawn_ray(si.p, si.n, ray.d); ray.o = r2.o; }
        ray.maxt = remaining_dist;
        needs_intersection |= active_surface;
        active &= (active_medium || active_surface) && any_nonzero(transmittance);
        if (active) total_dist += active_medium ? mei.t : si.t;
        if (active_surface && mesh_is_transition(sc, si.mesh)) medium = target_medium(sc, si, ray.d);
    }
    return emitter_val * transmittance;
}

// PRBVolpathIntegrator.sample (:96-351).  PRIMAL: returns L.  Adjoint: L starts at state_in and
// every contribution is subtracted as it is replayed.
template <bool PRIMAL>
LR_DEV LaneResult sample_prb(const DScene &sc, const DTri *tris_gid, Sampler &smp, Ray ray, V3 dL, V3 state_in, Counters &cnt) {
    uint32_t depth = 0; V3 L = PRIMAL ? mk3(0.f) : state_in, throughput = mk3(1.f); float eta = 1.f; bool active = true;
    SurfaceInteraction si = invalid_si(); bool needs_intersection = true; V3 last_scatter_p = mk3(0.f); float last_scatter_direction_pdf = 1.f;
    int medium = -1; bool valid_ray = false, specular_chain = true;
    uint32_t channel = (uint32_t) fminf(3.f * smp.next_1d(), 2.f);
    const float sgn = PRIMAL ? 1.f : -1.f;
    while (active) {
        active &= any_nonzero(throughput);
        float q = fminf(hmax(throughput) * sqr(eta), 0.99f);
        bool perform_rr = depth > (uint32_t) sc.rr_depth;
        if (active) { float u = smp.next_1d(); active &= (u < q) || !perform_rr; }
        if (perform_rr) throughput *= 1.f / q;
        if (!active) break;
        bool active_medium = medium >= 0, active_surface = !active_medium;
        MediumInteraction mei; mei.t = LR_INF; mei.medium = -1; mei.wi = mk3(0.f); mei.p = mk3(0.f);
        V3 weight = mk3(1.f), dweight = mk3(0.f); GridTap tap;
        bool escaped_medium = false, act_null_scatter = false, act_medium_scatter = false; float scatter_prob = 1.f;
        if (active_medium) {
            const DMedium &m = sc.media[medium];
            float u = smp.next_1d();
            mei = sample_interaction<!PRIMAL>(sc, medium, ray, u, channel, &tap, cnt.collisions);
            if (m.m.kind == LR_MEDIUM_HOMOGENEOUS && mei_valid(mei)) ray.maxt = mei.t;
            if (needs_intersection) si = intersect(sc, tris_gid, ray, cnt);
            needs_intersection = false;
            if (si.t < mei.t) mei.t = LR_INF;
            float t = fminf(mei.t, si.t) - mei.mint;
            V3 tr = exp3(mei.combined_extinction * (-t));
            V3 ffp = si.t < mei.t ? tr : tr * mei.combined_extinction;
            float tr_pdf = idx3(ffp, channel);
            weight *= tr_pdf > 0.f ? tr / tr_pdf : mk3(0.f);
            escaped_medium = !mei_valid(mei); active_medium = mei_valid(mei);
            if (sc.prb_handle_null) {
                if (active_medium) {
                    scatter_prob = hmean(mei.sigma_t / mei.combined_extinction);
                    act_null_scatter = smp.next_1d() >= scatter_prob; act_medium_scatter = !act_null_scatter;
                    if (act_null_scatter) { V3 A = weight; weight *= mei.sigma_n / (1.f - scatter_prob); dweight = A * (-m.m.scale / (1.f - scatter_prob)); }
                }
            } else act_medium_scatter = active_medium;
            if (act_medium_scatter) { depth += 1; last_scatter_p = mei.p; cnt.real_scatters++; }
        }
        active &= depth < sc.max_depth;
        act_medium_scatter &= active;
        if (act_null_scatter) { ray.o = mei.p; si.t = si.t - mei.t; }
        if (act_medium_scatter) {
            const DMedium &m = sc.media[mei.medium];
            V3 A = weight; weight *= mei.sigma_s / scatter_prob;
            dweight = A * medium_albedo(sc, m, mei.p) * (m.m.scale / scatter_prob);
            if (!PRIMAL) albedo_grad_scatter(sc, m, mei.p, albedo_grad_value(dL, L, weight, A * mei.sigma_t / scatter_prob));
        }
        throughput *= weight;
        if (!PRIMAL) {
            bool ev = (act_null_scatter || act_medium_scatter) && sc.media[mei.medium].m.kind == LR_MEDIUM_HETEROGENEOUS;
            if (ev) {
                const DGrid &g = sc.grids[sc.media[mei.medium].m.sigma_t_grid];
                float gsum = (dL.x * (L.x / fmaxf(1e-8f, weight.x))) * dweight.x + (dL.y * (L.y / fmaxf(1e-8f, weight.y))) * dweight.y +
                             (dL.z * (L.z / fmaxf(1e-8f, weight.z))) * dweight.z;
                if (g.differentiable && g.ch == 1 && gsum != 0.f) grid_grad_scatter(g, tap, gsum, true);
            }
        }
        Phase phase; phase.kind = LR_PHASE_ISOTROPIC; phase.g = 0.f;
        if (act_medium_scatter) { phase.kind = sc.media[mei.medium].m.phase; phase.g = sc.media[mei.medium].m.phase_g; }

        active_surface |= escaped_medium;
        bool intersect_now = active_surface && needs_intersection;
        if (intersect_now) si = intersect(sc, tris_gid, ray, cnt);
        if (sc.hide_emitters && intersect_now && depth == 0u && si_valid(si) && sc.meshes[si.mesh].emitter >= 0) {   // prbvolpath.py:218-231
            Ray r2 = spawn_ray(si.p, si.n, ray.d); Hit pi = skip_area_emitters(sc, tris_gid, r2, cnt);
            si = compute_si(sc, ray, pi, tris_gid);                       // the reference recomputes with the ORIGINAL ray here
        }
        if (active_surface) {
            bool count_direct = depth == 0u || specular_chain;
            int emitter = emitter_of(sc, si);
            if (emitter >= 0 && !(depth == 0u && sc.hide_emitters)) {
                DirectionSample ds = ds_from_si(sc, si, last_scatter_p);
                float emitter_pdf = pdf_emitter_direction(sc, ds);
                V3 emitted = eval_emitter_at(sc, emitter, si);
                V3 contrib = count_direct ? throughput * emitted : throughput * mis_weight(last_scatter_direction_pdf, emitter_pdf) * emitted;
                L += contrib * sgn;
            }
        }
        active_surface &= si_valid(si);
        {   // NEE (:259-291)
            bool active_e_surface = false;
            if (active_surface) { BSDF b(sc, si); active_e_surface = (b.flags() & F_SMOOTH) && (depth + 1 < sc.max_depth); }
            bool sample_emitters = act_medium_scatter ? sc.media[mei.medium].m.sample_emitters != 0 : false;
            if (act_medium_scatter) specular_chain = !sample_emitters;
            bool active_e_medium = act_medium_scatter && sample_emitters;
            if (active_e_surface || active_e_medium) {
                Sampler nee_sampler = smp; DirectionSample ds;
                V3 rp = active_e_medium ? mei.p : si.p, rn = active_e_medium ? mk3(0.f) : si.n;
                V3 emitted = sample_emitter_prb<false>(sc, tris_gid, rp, rn, active_e_surface ? &si : nullptr, smp, medium, channel, ds, mk3(0.f), mk3(0.f), cnt);
                V3 nee_weight; float nee_pdf;
                if (active_e_surface) { BSDF b(sc, si); nee_weight = b.eval_pdf(si.wi, si.sh.to_local(ds.d), nee_pdf); }
                else { float pv = phase.eval_pdf(mei.wi, ds.d); nee_weight = mk3(pv); nee_pdf = pv; }
                V3 contrib = throughput * nee_weight * mis_weight(ds.pdf, ds.delta ? 0.f : nee_pdf) * emitted;
                L += contrib * sgn;
                if (!PRIMAL) { DirectionSample ds2; sample_emitter_prb<true>(sc, tris_gid, rp, rn, active_e_surface ? &si : nullptr, nee_sampler, medium, channel, ds2, contrib, dL, cnt); }
            }
        }
        valid_ray |= act_medium_scatter;
        if (act_medium_scatter) {
            (void) smp.next_1d(); V2 s2 = smp.next_2d();
            float phase_pdf; V3 wo = phase.sample(mei.wi, s2, phase_pdf);
            act_medium_scatter &= phase_pdf > 0.f;
            if (act_medium_scatter) { ray.o = mei.p; ray.d = wo; ray.maxt = LR_LARGEST; needs_intersection = true; last_scatter_direction_pdf = phase_pdf; }
        }
        if (active_surface) {
            BSDF bsdf(sc, si);
            float s1 = smp.next_1d(); V2 s2 = smp.next_2d();
            V3 bsdf_weight; BSDFSample bs = bsdf.sample(si.wi, s1, s2, bsdf_weight);
            active_surface &= bs.pdf > 0.f;
            if (active_surface) {
                throughput *= bsdf_weight; eta *= bs.eta;
                ray = spawn_ray(si.p, si.n, si.sh.to_world(bs.wo)); needs_intersection = true;
                bool non_null = !(bs.sampled_type & F_NULL);
                if (non_null) { depth += 1; last_scatter_p = si.p; last_scatter_direction_pdf = bs.pdf; }
                valid_ray |= non_null;
                specular_chain |= non_null && (bs.sampled_type & F_DELTA);
                specular_chain &= !(bs.sampled_type & F_SMOOTH);
                if (mesh_is_transition(sc, si.mesh)) medium = target_medium(sc, si, ray.d);
            }
        }
        active &= (active_surface || active_medium);
    }
    LaneResult out; out.L = L; out.valid = valid_ray; return out;
}

} // namespace lr
