/* TEST INFRASTRUCTURE ONLY.  Included inside struct orc::Integrator.
 * volpath (src/integrators/volpath.cpp:93-554), one lane, JIT masking. */

/* VolumetricPathIntegrator::sample_emitter (volpath.cpp:400-554).
 * `ref_is_surface`: Interaction is a SurfaceInteraction (medium may change at
 * the boundary the reference point sits on, :419-421). */
V3 sample_emitter_vol(const Interaction &ref, const SurfaceInteraction *ref_si, Sampler &smp, int medium,
                      uint32_t channel, DirectionSample &ds_out) {
    V3 transmittance(1.f), emitter_val;
    DirectionSample ds = em.sample_emitter_direction(ref, smp.next_2d(), false, emitter_val);
    if (ds.pdf == 0.f) emitter_val = V3(0.f);
    bool active = ds.pdf != 0.f;
    ds_out = ds;
    if (!active) return emitter_val;

    Ray ray = ref.spawn_ray_to(ds.p);
    float max_dist = ray.maxt;
    if (ref_si && sc.meshes[ref_si->mesh].is_medium_transition()) medium = sc.target_medium(*ref_si, ray.d);

    float total_dist = 0.f; SurfaceInteraction si;   /* dr::zeros => t = inf (interaction.h:125-131) */
    bool needs_intersection = true;

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
                float tr_pdf = index_spectrum(ffp, channel);
                transmittance *= tr_pdf > 0.f ? tr / tr_pdf : V3(0.f);
            }
            if (mei.t > remaining_dist && mei.is_valid()) total_dist = ds.dist;
            if (mei.t > remaining_dist) mei.t = Inf;
            escaped_medium = !mei.is_valid();
            active_medium = mei.is_valid();
            is_spectral &= active_medium; not_spectral &= active_medium;
            if (active_medium) {
                total_dist += mei.t;
                ray.o = mei.p; si.t = si.t - mei.t;
                if (is_spectral) transmittance *= mei.sigma_n;
                if (not_spectral) transmittance *= mei.sigma_n / mei.combined_extinction;
            }
        }
        bool isect = active_surface && needs_intersection;
        if (isect) si = intersect(ray);
        if (isect) needs_intersection = false;
        active_surface |= escaped_medium;
        if (active_surface) total_dist += si.t;
        active_surface &= si.is_valid() && !active_medium;
        if (active_surface) {
            BSDF bsdf(sc, si);
            transmittance *= bsdf.eval_null_transmission();
        }
        if (active_surface) { Ray r2 = si.spawn_ray(ray.d); ray.o = r2.o; ray.d = r2.d; }
        ray.maxt = remaining_dist;
        needs_intersection |= active_surface;
        active &= (active_medium || active_surface) && any_nonzero(transmittance);
        if (active_surface && sc.meshes[si.mesh].is_medium_transition()) medium = sc.target_medium(si, ray.d);
    }
    return transmittance * emitter_val;
}

/* VolumetricPathIntegrator::sample (volpath.cpp:93-396) */
LaneResult sample_volpath(Sampler &smp, Ray ray, int initial_medium) {
    bool valid_ray = !hide_emitters && sc.env >= 0;
    float eta = 1.f; V3 throughput(1.f), result(0.f);
    int medium = initial_medium; MediumInteraction mei;
    bool specular_chain = !hide_emitters; uint32_t depth = 0;
    uint32_t channel = (uint32_t) std::min(smp.next_1d() * 3.f, 2.f);   /* :117-121 (n_channels - 1 as float min) */
    SurfaceInteraction si;
    bool needs_intersection = true;
    Interaction last_scatter_event;
    float last_scatter_direction_pdf = 1.f;
    bool active = true;
    while (active) {
        /* ---- termination / RR (:194-200) ---- */
        active &= any_nonzero(throughput);
        float q = std::min(hmax(throughput) * sqr(eta), .95f);
        bool perform_rr = depth > (uint32_t) rr_depth;
        if (active) { float u = smp.next_1d(); active &= (u < q) || !perform_rr; }
        if (perform_rr) throughput *= 1.f / q;
        active &= depth < max_depth;
        if (!active) break;

        bool active_medium = medium >= 0, active_surface = !active_medium;
        bool act_null_scatter = false, act_medium_scatter = false, escaped_medium = false;
        bool is_spectral = false, not_spectral = false;
        if (active_medium) { is_spectral = sc.media[medium].m.has_spectral_extinction != 0; not_spectral = !is_spectral; }

        if (active_medium) {
            const Medium &m = sc.media[medium];
            mei = sample_interaction(medium, ray, smp.next_1d(), channel);          /* :219 */
            if (m.m.kind == LR_MEDIUM_HOMOGENEOUS && mei.is_valid()) ray.maxt = mei.t;
            if (needs_intersection) si = intersect(ray);
            needs_intersection = false;
            if (si.t < mei.t) mei.t = Inf;
            if (is_spectral) {                                                        /* :228-232, medium.cpp:95-104 */
                float t = std::min(mei.t, si.t) - mei.mint;
                V3 tr = exp3(mei.combined_extinction * (-t));
                V3 pdf = si.t < mei.t ? tr : tr * mei.combined_extinction;
                float tr_pdf = index_spectrum(pdf, channel);
                throughput *= tr_pdf > 0.f ? tr / tr_pdf : V3(0.f);
            }
            escaped_medium = !mei.is_valid();
            active_medium = mei.is_valid();
            if (active_medium) {
                float null_scatter_prob = hmean(mei.sigma_n / mei.combined_extinction);
                bool null_scatter = smp.next_1d() < null_scatter_prob;             /* :239 */
                act_null_scatter = null_scatter; act_medium_scatter = !null_scatter;
                if (is_spectral && act_null_scatter) throughput *= mei.sigma_n / null_scatter_prob;
                if (act_medium_scatter) { depth += 1; last_scatter_event = mei; cnt.real_scatters++; }
            }
        }
        active &= depth < max_depth;                                                  /* :253 */
        act_medium_scatter &= active;
        if (act_null_scatter) { ray.o = mei.p; si.t = si.t - mei.t; }

        if (act_medium_scatter) {
            if (is_spectral) throughput *= mei.sigma_s / hmean(mei.sigma_t / mei.combined_extinction);
            if (not_spectral) throughput *= mei.sigma_s / mei.sigma_t;
            const Medium &m = sc.media[mei.medium]; Phase phase = phase_of(mei.medium);
            bool sample_emitters = m.m.sample_emitters != 0;
            valid_ray = true;
            specular_chain = false;
            specular_chain |= !sample_emitters;
            if (sample_emitters) {                                                    /* :276-283 */
                DirectionSample ds;
                V3 emitted = sample_emitter_vol(mei, nullptr, smp, medium, channel, ds);
                float phase_val = phase.eval_pdf(mei.wi, ds.d);
                result += throughput * phase_val * emitted * mis_weight(ds.pdf, ds.delta ? 0.f : phase_val);
            }
            float s1 = smp.next_1d(); (void) s1; V2 s2 = smp.next_2d();                /* :287-290 */
            float phase_pdf; V3 wo = phase.sample(mei.wi, s2, phase_pdf);
            if (phase_pdf > 0.f) {
                ray = mei.spawn_ray_m(wo);
                needs_intersection = true;
                last_scatter_direction_pdf = phase_pdf;
                /* phase_weight = 1 */
            }
        }

        /* ---- surfaces (:300-389) ---- */
        active_surface |= escaped_medium;
        bool intersect_now = active_surface && needs_intersection;
        if (intersect_now) si = intersect(ray);
        if (hide_emitters && intersect_now && depth == 0u && si.is_valid() && sc.meshes[si.mesh].emitter >= 0) {
            /* hide area emitters on the first segment (volpath.cpp:307-323: si recomputed with the skip ray) */
            Ray r2 = si.spawn_ray(ray.d); Hit pi = skip_area_emitters(r2);
            si = sc.compute_si(r2, pi);
        }
        if (active_surface) {
            bool count_direct = depth == 0u || specular_chain;
            int emitter = sc.emitter_of(si);
            bool active_e = emitter >= 0 && !(depth == 0u && hide_emitters);
            if (active_e) {
                float emitter_pdf = 1.f;
                if (!count_direct) {
                    DirectionSample ds = em.ds_from_si(si, last_scatter_event);
                    emitter_pdf = em.pdf_emitter_direction(last_scatter_event, ds);
                }
                V3 emitted = em.eval(emitter, si);
                result += count_direct ? throughput * emitted
                                       : throughput * mis_weight(last_scatter_direction_pdf, emitter_pdf) * emitted;
            }
        }
        active_surface &= si.is_valid();
        if (active_surface) {
            BSDF bsdf(sc, si);
            bool active_e = (bsdf.flags() & F_SMOOTH) && (depth + 1 < max_depth);
            if (active_e) {                                                            /* :351-363 */
                DirectionSample ds;
                V3 emitted = sample_emitter_vol(si, &si, smp, medium, channel, ds);
                V3 wo = si.to_local(ds.d);
                V3 bsdf_val = bsdf.eval(si.wi, wo); float bsdf_pdf = bsdf.pdf(si.wi, wo);
                result += throughput * bsdf_val * mis_weight(ds.pdf, ds.delta ? 0.f : bsdf_pdf) * emitted;
            }
            float s1 = smp.next_1d(); V2 s2 = smp.next_2d();                            /* :366-367 */
            V3 bsdf_val; BSDFSample bs = bsdf.sample(si.wi, s1, s2, bsdf_val);
            throughput *= bsdf_val; eta *= bs.eta;
            ray = si.spawn_ray(si.to_world(bs.wo));
            needs_intersection = true;
            bool non_null = !(bs.sampled_type & F_NULL);
            if (non_null) { depth += 1; last_scatter_event = si; last_scatter_direction_pdf = bs.pdf; }
            valid_ray |= non_null;
            specular_chain |= non_null && (bs.sampled_type & F_DELTA);
            specular_chain &= !(bs.sampled_type & F_SMOOTH);
            if (sc.meshes[si.mesh].is_medium_transition()) medium = sc.target_medium(si, ray.d);
        }
        active &= (active_surface || active_medium);
    }
    return {result, valid_ray};
}
