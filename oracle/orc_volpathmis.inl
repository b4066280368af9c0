/* TEST INFRASTRUCTURE ONLY.  Included inside struct orc::Integrator.
 * volpathmis (src/integrators/volpathmis.cpp:97-700), one lane, JIT masking.  Same draws and the same geometry as
 * volpath; what changes is the bookkeeping: instead of a throughput the path carries p / f per colour channel -- with
 * spectral MIS (use_spectral_mis, the default) a 3 x 3 matrix, row i = colour channel i, column j = the pdf the path would
 * have had with hero channel j -- and a second such quantity for the emitter-sampling strategy. */
struct MisW {
    V3 r[3]; bool spectral;
    static MisW ones(bool spectral) { MisW w; w.spectral = spectral; w.r[0] = w.r[1] = w.r[2] = V3(1.f); return w; }
};
static V3 fin_or_0(V3 a) { return V3(std::isfinite(a.x) ? a.x : 0.f, std::isfinite(a.y) ? a.y : 0.f, std::isfinite(a.z) ? a.z : 0.f); }
static V3 nan_to_0(V3 a) { return V3(std::isnan(a.x) ? 0.f : a.x, std::isnan(a.y) ? 0.f : a.y, std::isnan(a.z) ? 0.f : a.z); }
/* update_weights (volpathmis.cpp:636-653) */
static void update_weights(MisW &w, V3 p, V3 f, uint32_t channel, bool active) {
    if (w.spectral) {
        for (int i = 0; i < 3; ++i) {
            V3 ratio = p / f[i];                 /* Spectrum / Float = Spectrum * rcp(Float) */
            ratio = fin_or_0(ratio);
            ratio = ratio * w.r[i];
            if (active) w.r[i] = nan_to_0(ratio);
        }
    } else {
        float pdf = index_spectrum(p, channel);
        V3 ratio = w.r[0] * V3(pdf / f.x, pdf / f.y, pdf / f.z);
        if (active) w.r[0] = fin_or_0(ratio);
    }
}
static float hsum3(V3 a) { return (a.x + a.y) + a.z; }
static float hmin_abs(V3 a) { return std::min(fabsf(a.x), std::min(fabsf(a.y), fabsf(a.z))); }
/* mis_weight (volpathmis.cpp:655-669) */
static V3 mis_weight1(const MisW &w) {
    if (w.spectral) {
        V3 o;
        for (int i = 0; i < 3; ++i) { float s = hsum3(w.r[i]); o[i] = s == 0.f ? 0.f : 3.f / s; }
        return o;
    }
    if (hmin_abs(w.r[0]) == 0.f) return V3(0.f);
    return V3(1.f / w.r[0].x, 1.f / w.r[0].y, 1.f / w.r[0].z);
}
/* mis_weight (volpathmis.cpp:672-688) */
static V3 mis_weight2(const MisW &a, const MisW &b) {
    if (a.spectral) {
        V3 o;
        for (int i = 0; i < 3; ++i) { float s = hsum3(a.r[i] + b.r[i]); o[i] = s == 0.f ? 0.f : 3.f / s; }
        return o;
    }
    V3 s = a.r[0] + b.r[0];
    if (hmin_abs(s) == 0.f) return V3(0.f);
    return V3(1.f / s.x, 1.f / s.y, 1.f / s.z);
}

/* sample_emitter (volpathmis.cpp:466-633): returns emitter_val; p_over_f_nee / p_over_f_uni by reference */
V3 sample_emitter_mis(const Interaction &ref, const SurfaceInteraction *ref_si, Sampler &smp, int medium, const MisW &p_over_f,
                      uint32_t channel, MisW &p_over_f_nee, MisW &p_over_f_uni, DirectionSample &ds_out) {
    p_over_f_nee = p_over_f; p_over_f_uni = p_over_f;
    V3 weight;
    DirectionSample ds = em.sample_emitter_direction(ref, smp.next_2d(), false, weight);
    V3 emitter_val = weight * ds.pdf;
    if (ds.pdf == 0.f) emitter_val = V3(0.f);
    bool active = ds.pdf != 0.f;
    update_weights(p_over_f_nee, V3(ds.pdf), V3(1.f), channel, active);
    ds_out = ds;
    if (!active) return emitter_val;

    Ray ray = ref.spawn_ray_to(ds.p);
    float max_dist = ray.maxt;
    if (ref_si && sc.meshes[ref_si->mesh].is_medium_transition()) medium = sc.target_medium(*ref_si, ray.d);
    float total_dist = 0.f; SurfaceInteraction si; bool needs_intersection = true;
    while (active) {
        float remaining_dist = max_dist - total_dist;
        ray.maxt = remaining_dist;
        active &= remaining_dist > 0.f;
        if (!active) break;
        bool escaped_medium = false, active_medium = medium >= 0, active_surface = !active_medium;
        if (active_medium) {
            const Medium &m = sc.media[medium];
            MediumInteraction mei = sample_interaction(medium, ray, smp.next_1d(), channel);
            if (m.m.kind == LR_MEDIUM_HOMOGENEOUS && mei.is_valid()) ray.maxt = std::min(mei.t, remaining_dist);
            if (needs_intersection) si = intersect(ray);
            if (si.t < mei.t) mei.t = Inf;
            needs_intersection = false;
            bool is_spectral = m.m.has_spectral_extinction != 0, not_spectral = !is_spectral;
            if (is_spectral) {
                float t = std::min(remaining_dist, std::min(mei.t, si.t)) - mei.mint;
                V3 tr = exp3(mei.combined_extinction * (-t));
                V3 ffp = (si.t < mei.t || mei.t > remaining_dist) ? tr : tr * mei.combined_extinction;
                update_weights(p_over_f_nee, ffp, tr, channel, true);
                update_weights(p_over_f_uni, ffp, tr, channel, true);
            }
            if (mei.t > remaining_dist && mei.is_valid()) total_dist = ds.dist;
            if (mei.t > remaining_dist) mei.t = Inf;
            escaped_medium = !mei.is_valid();
            active_medium = mei.is_valid();
            is_spectral &= active_medium; not_spectral &= active_medium;
            if (active_medium) {
                total_dist += mei.t;
                ray.o = mei.p; si.t = si.t - mei.t;
                if (is_spectral) {
                    update_weights(p_over_f_nee, V3(1.f), mei.sigma_n, channel, true);
                    update_weights(p_over_f_uni, V3(hmean(mei.sigma_n / mei.combined_extinction)), mei.sigma_n, channel, true);
                }
                if (not_spectral) {
                    update_weights(p_over_f_nee, V3(1.f), mei.sigma_n / mei.combined_extinction, channel, true);
                    update_weights(p_over_f_uni, mei.sigma_n, mei.sigma_n, channel, true);
                }
            }
        }
        bool isect = active_surface && needs_intersection;
        if (isect) si = intersect(ray);
        active_surface |= escaped_medium;
        if (active_surface) total_dist += si.t;
        active_surface &= si.is_valid() && !active_medium;
        if (active_surface) {
            BSDF bsdf(sc, si);
            V3 bsdf_val = bsdf.eval_null_transmission();
            update_weights(p_over_f_nee, V3(1.f), bsdf_val, channel, true);
            update_weights(p_over_f_uni, V3(1.f), bsdf_val, channel, true);
        }
        if (active_surface) { Ray r2 = si.spawn_ray(ray.d); ray.o = r2.o; ray.d = r2.d; }
        ray.maxt = remaining_dist;
        needs_intersection |= active_surface;
        if (p_over_f.spectral) active &= (active_medium || active_surface) && any_nonzero(mis_weight1(p_over_f_uni));
        else active &= (active_medium || active_surface) && (any_nonzero(p_over_f_uni.r[0]) || any_nonzero(p_over_f_nee.r[0]));
        if (active_surface && sc.meshes[si.mesh].is_medium_transition()) medium = sc.target_medium(si, ray.d);
    }
    return emitter_val;
}

/* VolpathMisIntegratorImpl::sample (volpathmis.cpp:120-463) */
LaneResult sample_volpathmis(Sampler &smp, Ray ray, int initial_medium) {
    const bool spectral = sc.integ.spectral_mis != 0;
    bool valid_ray = !hide_emitters && sc.env >= 0;
    float eta = 1.f; V3 result(0.f);
    int medium = initial_medium; MediumInteraction mei;
    bool specular_chain = !hide_emitters; uint32_t depth = 0;
    MisW p_over_f = MisW::ones(spectral), p_over_f_nee = MisW::ones(spectral);
    uint32_t channel = (uint32_t) std::min(smp.next_1d() * 3.f, 2.f);
    SurfaceInteraction si;
    bool needs_intersection = true, last_event_was_null = false;
    Interaction last_scatter_event;
    bool active = true;
    while (active) {
        /* ---- termination (:222-234) ---- */
        V3 mis_throughput = mis_weight1(p_over_f);
        float q = std::min(hmax(mis_throughput) * sqr(eta), .95f);
        bool perform_rr = !last_event_was_null && depth > (uint32_t) rr_depth;
        { float u = smp.next_1d(); active &= !(u >= q && perform_rr); }
        update_weights(p_over_f, V3(q), V3(1.f), channel, perform_rr);
        last_event_was_null = false;
        active &= depth < max_depth;
        active &= any_nonzero(mis_weight1(p_over_f));
        if (!active) break;

        bool active_medium = medium >= 0, active_surface = !active_medium;
        bool act_null_scatter = false, act_medium_scatter = false, escaped_medium = false;
        bool is_spectral = false, not_spectral = false;
        if (active_medium) { is_spectral = sc.media[medium].m.has_spectral_extinction != 0; not_spectral = !is_spectral; }

        if (active_medium) {
            const Medium &m = sc.media[medium];
            mei = sample_interaction(medium, ray, smp.next_1d(), channel);          /* :256 */
            if (m.m.kind == LR_MEDIUM_HOMOGENEOUS && mei.is_valid()) ray.maxt = mei.t;
            if (needs_intersection) si = intersect(ray);
            needs_intersection = false;
            if (si.t < mei.t) mei.t = Inf;
            if (is_spectral) {                                                        /* :266-270 */
                float t = std::min(mei.t, si.t) - mei.mint;
                V3 tr = exp3(mei.combined_extinction * (-t));
                V3 pdf = si.t < mei.t ? tr : tr * mei.combined_extinction;
                update_weights(p_over_f, pdf, tr, channel, true);
                update_weights(p_over_f_nee, pdf, tr, channel, true);
            }
            escaped_medium = !mei.is_valid();
            active_medium = mei.is_valid();
            is_spectral &= active_medium; not_spectral &= active_medium;
        }
        if (active_medium) {                                                          /* :279-349 */
            float null_scatter_prob = hmean(mei.sigma_n / mei.combined_extinction);
            bool null_scatter = smp.next_1d() < null_scatter_prob;
            act_null_scatter = null_scatter; act_medium_scatter = !null_scatter;
            last_event_was_null = act_null_scatter;
            if (act_medium_scatter) { depth += 1; last_scatter_event = mei; cnt.real_scatters++; }
            bool sample_emitters = sc.media[mei.medium].m.sample_emitters != 0;
            active &= depth < max_depth;
            act_medium_scatter &= active;
            specular_chain &= !act_medium_scatter;
            specular_chain |= act_medium_scatter && !sample_emitters;
            if (act_null_scatter) {
                if (is_spectral) {
                    update_weights(p_over_f, V3(null_scatter_prob), mei.sigma_n, channel, true);
                    update_weights(p_over_f_nee, V3(1.f), mei.sigma_n, channel, true);
                }
                if (not_spectral) {
                    update_weights(p_over_f, mei.sigma_n, mei.sigma_n, channel, true);
                    update_weights(p_over_f_nee, V3(1.f), V3(null_scatter_prob), channel, true);
                }
                ray.o = mei.p; si.t = si.t - mei.t;
            }
            if (act_medium_scatter) {
                if (is_spectral) update_weights(p_over_f, V3(1.f - null_scatter_prob), mei.sigma_s, channel, true);
                if (not_spectral) update_weights(p_over_f, mei.sigma_t, mei.sigma_s, channel, true);
                Phase phase = phase_of(mei.medium);
                valid_ray = true;
                if (sample_emitters) {                                               /* :319-329 */
                    MisW nee_end, uni_end; DirectionSample ds;
                    V3 emitted = sample_emitter_mis(mei, nullptr, smp, medium, p_over_f, channel, nee_end, uni_end, ds);
                    bool active_e = true;
                    float phase_val = phase.eval_pdf(mei.wi, ds.d);
                    update_weights(nee_end, V3(1.f), V3(phase_val), channel, active_e);
                    update_weights(uni_end, V3(ds.delta ? 0.f : phase_val), V3(phase_val), channel, active_e);
                    result += mis_weight2(nee_end, uni_end) * emitted;
                }
                p_over_f_nee = p_over_f;
                float s1 = smp.next_1d(); (void) s1; V2 s2 = smp.next_2d();
                float phase_pdf; V3 wo = phase.sample(mei.wi, s2, phase_pdf);
                ray = mei.spawn_ray_m(wo);
                needs_intersection = true;
                update_weights(p_over_f, V3(phase_pdf), V3(1.f * phase_pdf), channel, true);
                update_weights(p_over_f_nee, V3(1.f), V3(1.f * phase_pdf), channel, true);
            }
        }

        /* ---- surfaces (:352-446) ---- */
        active_surface |= escaped_medium;
        bool intersect_now = active_surface && needs_intersection;
        if (intersect_now) si = intersect(ray);
        if (active_surface) {
            if (hide_emitters && depth == 0u && si.is_valid() && sc.meshes[si.mesh].emitter >= 0 && intersect_now) {   /* :360-375 */
                Ray r2 = si.spawn_ray(ray.d); Hit pi = skip_area_emitters(r2);
                si = sc.compute_si(r2, pi);
            }
            bool count_direct = depth == 0u || specular_chain;
            int emitter = sc.emitter_of(si);
            bool active_e = emitter >= 0 && !(depth == 0u && hide_emitters);
            if (active_e) {
                {   /* :387-392: the block is entered if ANY lane needs it; the update itself is masked by active_e only */
                    DirectionSample ds = em.ds_from_si(si, last_scatter_event);
                    float emitter_pdf = em.pdf_emitter_direction(last_scatter_event, ds);
                    update_weights(p_over_f_nee, V3(emitter_pdf), V3(1.f), channel, true);
                }
                V3 emitted = em.eval(emitter, si);
                result += count_direct ? mis_weight1(p_over_f) * emitted : mis_weight2(p_over_f, p_over_f_nee) * emitted;
            }
        }
        active_surface &= si.is_valid();
        if (active_surface) {
            BSDF bsdf(sc, si);
            bool active_e = (bsdf.flags() & F_SMOOTH) && (depth + 1 < max_depth);
            if (active_e) {                                                            /* :408-416 */
                MisW nee_end, uni_end; DirectionSample ds;
                V3 emitted = sample_emitter_mis(si, &si, smp, medium, p_over_f, channel, nee_end, uni_end, ds);
                V3 wo = si.to_local(ds.d);
                V3 bsdf_val = bsdf.eval(si.wi, wo); float bsdf_pdf = bsdf.pdf(si.wi, wo);
                update_weights(nee_end, V3(1.f), bsdf_val, channel, true);
                update_weights(uni_end, V3(ds.delta ? 0.f : bsdf_pdf), bsdf_val, channel, true);
                result += mis_weight2(nee_end, uni_end) * emitted;
            }
            float s1 = smp.next_1d(); V2 s2 = smp.next_2d();
            V3 bsdf_weight; BSDFSample bs = bsdf.sample(si.wi, s1, s2, bsdf_weight);
            bool invalid_bsdf_sample = bs.pdf == 0.f;
            active_surface &= bs.pdf > 0.f;
            if (active_surface) {
                eta *= bs.eta;
                ray = si.spawn_ray(si.to_world(bs.wo));
                needs_intersection = true;
            }
            bool non_null = active_surface && !(bs.sampled_type & F_NULL);
            valid_ray |= non_null || invalid_bsdf_sample;
            specular_chain |= non_null && (bs.sampled_type & F_DELTA);
            specular_chain &= !(active_surface && (bs.sampled_type & F_SMOOTH));
            if (non_null) { depth += 1; last_scatter_event = si; p_over_f_nee = p_over_f; }
            update_weights(p_over_f, V3(bs.pdf), bsdf_weight * bs.pdf, channel, active_surface);
            update_weights(p_over_f_nee, V3(1.f), bsdf_weight * bs.pdf, channel, non_null);
            if (active_surface && sc.meshes[si.mesh].is_medium_transition()) medium = sc.target_medium(si, ray.d);
        }
        active &= (active_surface || active_medium);
    }
    return {result, valid_ray};
}
