// lr_volpathmis.cuh -- volpathmis (src/integrators/volpathmis.cpp:97-700), device code, one lane per thread.
// Same draws and the same geometry as volpath (lr_integrators.cuh); the path carries p / f instead of a throughput: with spectral
// MIS (use_spectral_mis, the default) a 3 x 3 matrix -- row i = colour channel i, column j = the pdf the path would have had
// with hero channel j -- and a second one for the emitter-sampling strategy; without it one row.  SPECTRAL is a template
// parameter so that the one-row variant keeps 6 registers of weights instead of 18.
#pragma once
#include "lr_integrators.cuh"

namespace lr {

template <bool SPECTRAL> struct MisW { V3 r[SPECTRAL ? 3 : 1]; };
template <bool SPECTRAL> LR_DEV MisW<SPECTRAL> misw_ones() {
    MisW<SPECTRAL> w;
#pragma unroll
    for (int i = 0; i < (SPECTRAL ? 3 : 1); ++i) w.r[i] = mk3(1.f);
    return w;
}
LR_DEV V3 fin_or_0(V3 a) { return mk3(isfinite(a.x) ? a.x : 0.f, isfinite(a.y) ? a.y : 0.f, isfinite(a.z) ? a.z : 0.f); }
LR_DEV V3 nan_to_0(V3 a) { return mk3(isnan(a.x) ? 0.f : a.x, isnan(a.y) ? 0.f : a.y, isnan(a.z) ? 0.f : a.z); }
LR_DEV float v3_at(V3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

// update_weights (volpathmis.cpp:636-653)
template <bool SPECTRAL>
LR_DEV void update_weights(MisW<SPECTRAL> &w, V3 p, V3 f, uint32_t channel, bool active) {
    if (SPECTRAL) {
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            V3 ratio = p / v3_at(f, i);                       // Spectrum / Float = Spectrum * rcp(Float)
            ratio = fin_or_0(ratio);
            ratio = mk3(__fmul_rn(ratio.x, w.r[SPECTRAL ? i : 0].x), __fmul_rn(ratio.y, w.r[SPECTRAL ? i : 0].y), __fmul_rn(ratio.z, w.r[SPECTRAL ? i : 0].z));
            if (active) w.r[SPECTRAL ? i : 0] = nan_to_0(ratio);
        }
    } else {
        const float pdf = idx3(p, channel);
        V3 ratio = mk3(__fmul_rn(w.r[0].x, pdf / f.x), __fmul_rn(w.r[0].y, pdf / f.y), __fmul_rn(w.r[0].z, pdf / f.z));
        if (active) w.r[0] = fin_or_0(ratio);
    }
}
LR_DEV float hsum3(V3 a) { return __fadd_rn(__fadd_rn(a.x, a.y), a.z); }
LR_DEV float hmin_abs(V3 a) { return fminf(fabsf(a.x), fminf(fabsf(a.y), fabsf(a.z))); }
// mis_weight (volpathmis.cpp:655-669)
template <bool SPECTRAL>
LR_DEV V3 mis_weight1(const MisW<SPECTRAL> &w) {
    if (SPECTRAL) {
        float o[3];
#pragma unroll
        for (int i = 0; i < 3; ++i) { const float s = hsum3(w.r[SPECTRAL ? i : 0]); o[i] = s == 0.f ? 0.f : 3.f / s; }
        return mk3(o[0], o[1], o[2]);
    }
    if (hmin_abs(w.r[0]) == 0.f) return mk3(0.f);
    return mk3(1.f / w.r[0].x, 1.f / w.r[0].y, 1.f / w.r[0].z);
}
// mis_weight (volpathmis.cpp:672-688)
template <bool SPECTRAL>
LR_DEV V3 mis_weight2(const MisW<SPECTRAL> &a, const MisW<SPECTRAL> &b) {
    if (SPECTRAL) {
        float o[3];
#pragma unroll
        for (int i = 0; i < 3; ++i) { const float s = hsum3(a.r[SPECTRAL ? i : 0] + b.r[SPECTRAL ? i : 0]); o[i] = s == 0.f ? 0.f : 3.f / s; }
        return mk3(o[0], o[1], o[2]);
    }
    const V3 s = a.r[0] + b.r[0];
    if (hmin_abs(s) == 0.f) return mk3(0.f);
    return mk3(1.f / s.x, 1.f / s.y, 1.f / s.z);
}

// sample_emitter (volpathmis.cpp:466-633): returns emitter_val; p_over_f_nee / p_over_f_uni by reference
template <bool SPECTRAL>
LR_DEV V3 sample_emitter_mis(const DScene &sc, const DTri *tris_gid, V3 ref_p, V3 ref_n, const SurfaceInteraction *ref_si, Sampler &smp, int medium,
                             const MisW<SPECTRAL> &p_over_f, uint32_t channel, MisW<SPECTRAL> &p_over_f_nee, MisW<SPECTRAL> &p_over_f_uni,
                             DirectionSample &ds, Counters &cnt) {
    p_over_f_nee = p_over_f; p_over_f_uni = p_over_f;
    V3 weight;
    ds = sample_emitter_direction(sc, ref_p, smp.next_2d(), weight);
    V3 emitter_val = weight * ds.pdf;
    if (ds.pdf == 0.f) emitter_val = mk3(0.f);
    bool active = ds.pdf != 0.f;
    update_weights<SPECTRAL>(p_over_f_nee, mk3(ds.pdf), mk3(1.f), channel, active);
    if (!active) return emitter_val;
    Ray ray = spawn_ray_to(ref_p, ref_n, ds.p);
    const float max_dist = ray.maxt;
    if (ref_si && mesh_is_transition(sc, ref_si->mesh)) medium = target_medium(sc, *ref_si, ray.d);
    float total_dist = 0.f; SurfaceInteraction si = invalid_si(); bool needs_intersection = true;
    while (active) {
        const float remaining_dist = max_dist - total_dist;
        ray.maxt = remaining_dist;
        active &= remaining_dist > 0.f;
        if (!active) break;
        bool escaped_medium = false, active_medium = medium >= 0, active_surface = !active_medium;
        if (active_medium) {
            const DMedium &m = sc.media[medium];
            MediumInteraction mei = sample_interaction<false>(sc, medium, ray, smp.next_1d(), channel, nullptr, cnt.collisions);
            if (m.m.kind == LR_MEDIUM_HOMOGENEOUS && mei_valid(mei)) ray.maxt = fminf(mei.t, remaining_dist);
            if (needs_intersection) si = intersect(sc, tris_gid, ray, cnt);
            if (si.t < mei.t) mei.t = LR_INF;
            needs_intersection = false;
            bool is_spectral = m.m.has_spectral_extinction != 0, not_spectral = !is_spectral;
            if (is_spectral) {
                const float t = fminf(remaining_dist, fminf(mei.t, si.t)) - mei.mint;
                const V3 tr = exp3(mei.combined_extinction * (-t));
                const V3 ffp = (si.t < mei.t || mei.t > remaining_dist) ? tr : tr * mei.combined_extinction;
                update_weights<SPECTRAL>(p_over_f_nee, ffp, tr, channel, true);
                update_weights<SPECTRAL>(p_over_f_uni, ffp, tr, channel, true);
            }
            if (mei.t > remaining_dist && mei_valid(mei)) total_dist = ds.dist;
            if (mei.t > remaining_dist) mei.t = LR_INF;
            escaped_medium = !mei_valid(mei); active_medium = mei_valid(mei);
            is_spectral &= active_medium; not_spectral &= active_medium;
            if (active_medium) {
                total_dist += mei.t; ray.o = mei.p; si.t = si.t - mei.t;
                if (is_spectral) {
                    update_weights<SPECTRAL>(p_over_f_nee, mk3(1.f), mei.sigma_n, channel, true);
                    update_weights<SPECTRAL>(p_over_f_uni, mk3(hmean(mei.sigma_n / mei.combined_extinction)), mei.sigma_n, channel, true);
                }
                if (not_spectral) {
                    update_weights<SPECTRAL>(p_over_f_nee, mk3(1.f), mei.sigma_n / mei.combined_extinction, channel, true);
                    update_weights<SPECTRAL>(p_over_f_uni, mei.sigma_n, mei.sigma_n, channel, true);
                }
            }
        }
        if (active_surface && needs_intersection) si = intersect(sc, tris_gid, ray, cnt);
        active_surface |= escaped_medium;
        if (active_surface) total_dist += si.t;
        active_surface &= si_valid(si) && !active_medium;
        if (active_surface) {
            BSDF bsdf(sc, si);
            const V3 bsdf_val = mk3(bsdf.eval_null_transmission());
            update_weights<SPECTRAL>(p_over_f_nee, mk3(1.f), bsdf_val, channel, true);
            update_weights<SPECTRAL>(p_over_f_uni, mk3(1.f), bsdf_val, channel, true);
            Ray r2 = spawn_ray(si.p, si.n, ray.d); ray.o = r2.o;
        }
        ray.maxt = remaining_dist;
        needs_intersection |= active_surface;
        if (SPECTRAL) active &= (active_medium || active_surface) && any_nonzero(mis_weight1<SPECTRAL>(p_over_f_uni));
        else active &= (active_medium || active_surface) && (any_nonzero(p_over_f_uni.r[0]) || any_nonzero(p_over_f_nee.r[0]));
        if (active_surface && mesh_is_transition(sc, si.mesh)) medium = target_medium(sc, si, ray.d);
    }
    return emitter_val;
}

// VolpathMisIntegratorImpl::sample (volpathmis.cpp:120-463) as a resumable path: init() + one step() per iteration of the reference's
// loop.  This per-lane form is the parity probe; production is the wavefront (lr_wavefront_mis.cuh).  ncu on k_render<volpathmis>: 2.4 of 32
// threads active per instruction (profiles/r2_kernel_metrics_mis.txt) -- a step contains the emitter-sampling walk, tens of global-majorant
// collisions for one lane while its neighbours wait -- and refilling finished lanes between steps does not help (1.04x).
template <bool SPECTRAL>
struct MisPath {
    Sampler smp; Ray ray; SurfaceInteraction si; V3 result, last_scatter_p; float eta;
    MisW<SPECTRAL> p_over_f, p_over_f_nee;
    int medium; uint32_t depth, channel; bool valid_ray, specular_chain, needs_intersection, last_event_was_null;

    LR_DEV void init(const DScene &sc, const Sampler &s, const Ray &r, int initial_medium) {
        smp = s; ray = r; valid_ray = !sc.hide_emitters && sc.env >= 0;
        eta = 1.f; result = mk3(0.f); medium = initial_medium;
        specular_chain = !sc.hide_emitters; depth = 0;
        p_over_f = misw_ones<SPECTRAL>(); p_over_f_nee = misw_ones<SPECTRAL>();
        channel = (uint32_t) fminf(smp.next_1d() * 3.f, 2.f);
        si = invalid_si(); needs_intersection = true; last_event_was_null = false; last_scatter_p = mk3(0.f);
    }
    LR_DEV LaneResult finish() const { LaneResult out; out.L = result; out.valid = valid_ray; return out; }

    LR_DEV bool step(const DScene &sc, const DTri *tris_gid, Counters &cnt) {
        bool active = true;
        const V3 mis_throughput = mis_weight1<SPECTRAL>(p_over_f);                 // :222-234
        const float q = fminf(hmax(mis_throughput) * sqr(eta), .95f);
        const bool perform_rr = !last_event_was_null && depth > (uint32_t) sc.rr_depth;
        { const float u = smp.next_1d(); active &= !(u >= q && perform_rr); }
        update_weights<SPECTRAL>(p_over_f, mk3(q), mk3(1.f), channel, perform_rr);
        last_event_was_null = false;
        active &= depth < sc.max_depth;
        active &= any_nonzero(mis_weight1<SPECTRAL>(p_over_f));
        if (!active) return false;

        MediumInteraction mei; mei.t = LR_INF; mei.medium = -1;
        bool active_medium = medium >= 0, active_surface = !active_medium;
        bool act_null_scatter = false, act_medium_scatter = false, escaped_medium = false, is_spectral = false, not_spectral = false;
        if (active_medium) {
            const DMedium &m = sc.media[medium];
            is_spectral = m.m.has_spectral_extinction != 0; not_spectral = !is_spectral;
            mei = sample_interaction<false>(sc, medium, ray, smp.next_1d(), channel, nullptr, cnt.collisions);   // :256
            if (m.m.kind == LR_MEDIUM_HOMOGENEOUS && mei_valid(mei)) ray.maxt = mei.t;
            if (needs_intersection) si = intersect(sc, tris_gid, ray, cnt);
            needs_intersection = false;
            if (si.t < mei.t) mei.t = LR_INF;
            if (is_spectral) {                                                      // :266-270
                const float t = fminf(mei.t, si.t) - mei.mint;
                const V3 tr = exp3(mei.combined_extinction * (-t));
                const V3 pdf = si.t < mei.t ? tr : tr * mei.combined_extinction;
                update_weights<SPECTRAL>(p_over_f, pdf, tr, channel, true);
                update_weights<SPECTRAL>(p_over_f_nee, pdf, tr, channel, true);
            }
            escaped_medium = !mei_valid(mei); active_medium = mei_valid(mei);
            is_spectral &= active_medium; not_spectral &= active_medium;
        }
        if (active_medium) {                                                        // :279-349
            const float null_scatter_prob = hmean(mei.sigma_n / mei.combined_extinction);
            const bool null_scatter = smp.next_1d() < null_scatter_prob;
            act_null_scatter = null_scatter; act_medium_scatter = !null_scatter;
            last_event_was_null = act_null_scatter;
            if (act_medium_scatter) { depth += 1; last_scatter_p = mei.p; cnt.real_scatters++; }
            const DMedium &m = sc.media[mei.medium];
            const bool sample_emitters = m.m.sample_emitters != 0;
            active &= depth < sc.max_depth;
            act_medium_scatter &= active;
            specular_chain &= !act_medium_scatter;
            specular_chain |= act_medium_scatter && !sample_emitters;
            if (act_null_scatter) {
                if (is_spectral) {
                    update_weights<SPECTRAL>(p_over_f, mk3(null_scatter_prob), mei.sigma_n, channel, true);
                    update_weights<SPECTRAL>(p_over_f_nee, mk3(1.f), mei.sigma_n, channel, true);
                }
                if (not_spectral) {
                    update_weights<SPECTRAL>(p_over_f, mei.sigma_n, mei.sigma_n, channel, true);
                    update_weights<SPECTRAL>(p_over_f_nee, mk3(1.f), mk3(null_scatter_prob), channel, true);
                }
                ray.o = mei.p; si.t = si.t - mei.t;
            }
            if (act_medium_scatter) {
                if (is_spectral) update_weights<SPECTRAL>(p_over_f, mk3(1.f - null_scatter_prob), mei.sigma_s, channel, true);
                if (not_spectral) update_weights<SPECTRAL>(p_over_f, mei.sigma_t, mei.sigma_s, channel, true);
                Phase phase; phase.kind = m.m.phase; phase.g = m.m.phase_g;
                valid_ray = true;
                if (sample_emitters) {                                             // :319-329
                    MisW<SPECTRAL> nee_end, uni_end; DirectionSample ds;
                    const V3 emitted = sample_emitter_mis<SPECTRAL>(sc, tris_gid, mei.p, mk3(0.f), nullptr, smp, medium, p_over_f, channel, nee_end, uni_end, ds, cnt);
                    const float phase_val = phase.eval_pdf(mei.wi, ds.d);
                    update_weights<SPECTRAL>(nee_end, mk3(1.f), mk3(phase_val), channel, true);
                    update_weights<SPECTRAL>(uni_end, mk3(ds.delta ? 0.f : phase_val), mk3(phase_val), channel, true);
                    result += mis_weight2<SPECTRAL>(nee_end, uni_end) * emitted;
                }
                p_over_f_nee = p_over_f;
                (void) smp.next_1d(); const V2 s2 = smp.next_2d();
                float phase_pdf; const V3 wo = phase.sample(mei.wi, s2, phase_pdf);
                ray.o = mei.p; ray.d = wo; ray.maxt = LR_LARGEST;
                needs_intersection = true;
                update_weights<SPECTRAL>(p_over_f, mk3(phase_pdf), mk3(1.f * phase_pdf), channel, true);
                update_weights<SPECTRAL>(p_over_f_nee, mk3(1.f), mk3(1.f * phase_pdf), channel, true);
            }
        }

        active_surface |= escaped_medium;                                           // :352-446
        const bool intersect_now = active_surface && needs_intersection;
        if (intersect_now) si = intersect(sc, tris_gid, ray, cnt);
        if (active_surface) {
            if (sc.hide_emitters && depth == 0u && si_valid(si) && sc.meshes[si.mesh].emitter >= 0 && intersect_now) {   // :360-375
                Ray r2 = spawn_ray(si.p, si.n, ray.d); Hit pi = skip_area_emitters(sc, tris_gid, r2, cnt);
                si = compute_si(sc, r2, pi, tris_gid);
            }
            const bool count_direct = depth == 0u || specular_chain;
            const int emitter = emitter_of(sc, si);
            if (emitter >= 0 && !(depth == 0u && sc.hide_emitters)) {
                {   // :387-392: the block is entered if ANY lane needs it; the update itself is masked by active_e only
                    DirectionSample ds = ds_from_si(sc, si, last_scatter_p);
                    const float emitter_pdf = pdf_emitter_direction(sc, ds);
                    update_weights<SPECTRAL>(p_over_f_nee, mk3(emitter_pdf), mk3(1.f), channel, true);
                }
                const V3 emitted = eval_emitter_at(sc, emitter, si);
                result += count_direct ? mis_weight1<SPECTRAL>(p_over_f) * emitted : mis_weight2<SPECTRAL>(p_over_f, p_over_f_nee) * emitted;
            }
        }
        active_surface &= si_valid(si);
        if (active_surface) {
            BSDF bsdf(sc, si);
            if ((bsdf.flags() & F_SMOOTH) && (depth + 1 < sc.max_depth)) {          // :408-416
                MisW<SPECTRAL> nee_end, uni_end; DirectionSample ds;
                const V3 emitted = sample_emitter_mis<SPECTRAL>(sc, tris_gid, si.p, si.n, &si, smp, medium, p_over_f, channel, nee_end, uni_end, ds, cnt);
                const V3 wo = si.sh.to_local(ds.d);
                float bsdf_pdf; const V3 bsdf_val = bsdf.eval_pdf(si.wi, wo, bsdf_pdf);
                update_weights<SPECTRAL>(nee_end, mk3(1.f), bsdf_val, channel, true);
                update_weights<SPECTRAL>(uni_end, mk3(ds.delta ? 0.f : bsdf_pdf), bsdf_val, channel, true);
                result += mis_weight2<SPECTRAL>(nee_end, uni_end) * emitted;
            }
            const float s1 = smp.next_1d(); const V2 s2 = smp.next_2d();
            V3 bsdf_weight; BSDFSample bs = bsdf.sample(si.wi, s1, s2, bsdf_weight);
            const bool invalid_bsdf_sample = bs.pdf == 0.f;
            active_surface &= bs.pdf > 0.f;
            const bool non_null = active_surface && !(bs.sampled_type & F_NULL);
            valid_ray |= non_null || invalid_bsdf_sample;
            specular_chain |= non_null && (bs.sampled_type & F_DELTA);
            specular_chain &= !(active_surface && (bs.sampled_type & F_SMOOTH));
            if (non_null) { depth += 1; last_scatter_p = si.p; p_over_f_nee = p_over_f; }
            update_weights<SPECTRAL>(p_over_f, mk3(bs.pdf), bsdf_weight * bs.pdf, channel, active_surface);
            update_weights<SPECTRAL>(p_over_f_nee, mk3(1.f), bsdf_weight * bs.pdf, channel, non_null);
            if (active_surface) {
                eta *= bs.eta;
                ray = spawn_ray(si.p, si.n, si.sh.to_world(bs.wo));
                needs_intersection = true;
                if (mesh_is_transition(sc, si.mesh)) medium = target_medium(sc, si, ray.d);
            }
        }
        active &= (active_surface || active_medium);
        return active;
    }
};

template <bool SPECTRAL>
LR_DEV LaneResult sample_volpathmis(const DScene &sc, const DTri *tris_gid, Sampler &smp, Ray ray, int initial_medium, Counters &cnt) {
    MisPath<SPECTRAL> p; p.init(sc, smp, ray, initial_medium);
    while (p.step(sc, tris_gid, cnt)) { }
    smp = p.smp;
    return p.finish();
}

} // namespace lr
