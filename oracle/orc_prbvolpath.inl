/* TEST INFRASTRUCTURE ONLY.  Included inside struct orc::Integrator.
 * prbvolpath (src/python/python/ad/integrators/prbvolpath.py:96-444), one lane.
 * Backward mode accumulates d(loss)/d(sigma_t grid) with the closed-form
 * per-event derivative of the attached `weight` / `tr_multiplier`
 * (prbvolpath.py:169-204, :395-427): the only dependence on the grid is
 * sigma_t = scale*grid(p), sigma_n = majorant - sigma_t (majorant detached,
 * src/media/heterogeneous.cpp:163), sigma_s = sigma_t*albedo. */

bool prb_handle_null = false;       /* prepare_scene (:80-93): some medium is heterogeneous */
bool prb_nee_homogeneous = false;   /* some medium is homogeneous                            */
void prb_prepare() {
    for (const Mesh &me : sc.meshes)
        for (int mid : {me.interior, me.exterior})
            if (mid >= 0) {
                if (sc.media[mid].m.kind == LR_MEDIUM_HOMOGENEOUS) prb_nee_homogeneous = true; else prb_handle_null = true;
            }
}

/* scatter-add of `g` (already multiplied by dL/d sigma_t-grid-value) into the 8 corners */
void grid_grad_add(const MediumInteraction &mei, double g) {
    const Medium &m = sc.media[mei.medium];
    if (m.m.kind != LR_MEDIUM_HETEROGENEOUS) return;
    Grid &gr = const_cast<Grid &>(sc.grids[m.m.sigma_t_grid]);
    if (!gr.differentiable || gr.ch != 1) return;
    for (int k = 0; k < 8; ++k) {
        double v = g * (double) mei.cw[k];
        std::atomic_ref<double> a(gr.grad[mei.cidx[k]]);
        double old = a.load(std::memory_order_relaxed);
        while (!a.compare_exchange_weak(old, old + v, std::memory_order_relaxed)) {}
    }
}

/* gradient w.r.t. an ALBEDO grid value: sigma_s = sigma_t * albedo (heterogeneous.cpp:189-195), so at a real scatter
   d weight_c / d albedo_c = A_c * sigma_t_c / scatter_prob (prbvolpath.py:202-204 back-propagated through `mei.sigma_s`).
   `g`: per-channel dL * (L / weight) * d weight / d albedo. */
void albedo_grad_add(const MediumInteraction &mei, V3 g) {
    const Medium &m = sc.media[mei.medium];
    if (m.m.albedo_grid < 0) return;
    Grid &gr = const_cast<Grid &>(sc.grids[m.m.albedo_grid]);
    if (!gr.differentiable) return;
    float o[3]; uint32_t cidx[8]; float cw[8]; gr.eval(mei.p, o, cidx, cw);
    for (int k = 0; k < 8; ++k)
        for (uint32_t c = 0; c < 3; ++c) {
            double v = (double) g[c] * (double) cw[k];
            std::atomic_ref<double> a(gr.grad[(size_t) cidx[k] * gr.ch + (gr.ch == 1 ? 0 : c)]);
            double old = a.load(std::memory_order_relaxed);
            while (!a.compare_exchange_weak(old, old + v, std::memory_order_relaxed)) {}
        }
}

/* PRBVolpathIntegrator.sample_emitter (:354-444).  adj != nullptr => adjoint
 * replay with adj_emitted = *adj and the lane's dL. */
V3 sample_emitter_prb(const Interaction &ref, const SurfaceInteraction *ref_si, Sampler &smp, int medium, uint32_t channel,
                      DirectionSample &ds_out, const V3 *adj, V3 dL) {
    V3 emitter_val;
    DirectionSample ds = em.sample_emitter_direction(ref, smp.next_2d(), false, emitter_val);
    bool invalid = ds.pdf == 0.f; if (invalid) emitter_val = V3(0.f);
    bool active = !invalid; ds_out = ds;
    if (!active) medium = -1;
    if (ref_si && sc.meshes[ref_si->mesh].is_medium_transition()) medium = sc.target_medium(*ref_si, ds.d);
    if (!active) return V3(0.f);
    Ray ray = ref.spawn_ray_to(ds.p);
    float max_dist = ray.maxt, total_dist = 0.f; SurfaceInteraction si; bool needs_intersection = true; V3 transmittance(1.f);
    while (active) {
        float remaining_dist = max_dist - total_dist; ray.maxt = remaining_dist;
        active &= remaining_dist > 0.f;
        if (!active) break;
        if (needs_intersection) si = intersect(ray);
        needs_intersection = false;
        bool active_medium = medium >= 0, active_surface = !active_medium;
        MediumInteraction mei; V3 tr_multiplier(1.f); bool escaped_medium = false;
        if (active_medium) {
            const Medium &m = sc.media[medium];
            mei = sample_interaction(medium, ray, smp.next_1d(), channel);
            if (si.t < mei.t) mei.t = Inf;
            if (prb_nee_homogeneous && m.m.kind == LR_MEDIUM_HOMOGENEOUS) {
                float t = std::min(remaining_dist, si.t) - mei.mint;         /* transmittance_eval_pdf()[0] with mei.t = min(remaining, si.t) */
                tr_multiplier = exp3(mei.combined_extinction * (-t));
                mei.t = Inf;
            }
            escaped_medium = !mei.is_valid();
            active_medium = mei.is_valid();
            if (active_medium) {
                ray.o = mei.p; si.t = si.t - mei.t;
                tr_multiplier *= mei.sigma_n / mei.combined_extinction;
            }
        }
        active_surface |= escaped_medium;
        active_surface &= si.is_valid() && !active_medium;
        if (active_surface) { BSDF bsdf(sc, si); tr_multiplier *= bsdf.eval_null_transmission(); }
        if (adj && active_medium && sc.media[medium].m.kind == LR_MEDIUM_HETEROGENEOUS) {
            /* d tr_multiplier_c / d grid = -scale / majorant_c ; backward(tr_mult * dL*adj/tr_mult) */
            const Medium &m = sc.media[medium]; double g = 0.0;
            for (int c = 0; c < 3; ++c)
                if (tr_multiplier[c] > 0.f)
                    g += (double) (dL[c] * (*adj)[c] / tr_multiplier[c]) * (double) (-m.m.scale / mei.combined_extinction[c]);
            grid_grad_add(mei, g);
        }
        transmittance *= tr_multiplier;
        if (active_surface) { Ray r2 = si.spawn_ray(ray.d); ray.o = r2.o; ray.d = r2.d; }
        ray.maxt = remaining_dist;
        needs_intersection |= active_surface;
        active &= (active_medium || active_surface) && any_nonzero(transmittance);
        if (active) total_dist += active_medium ? mei.t : si.t;
        if (active_surface && sc.meshes[si.mesh].is_medium_transition()) medium = sc.target_medium(si, ray.d);
    }
    return emitter_val * transmittance;
}

/* PRBVolpathIntegrator.sample (:96-351).  primal: returns L.  backward: L starts
 * at state_in, contributions are subtracted as replayed, gradients scattered. */
LaneResult sample_prb(bool primal, Sampler &smp, Ray ray, V3 dL, V3 state_in) {
    uint32_t depth = 0; V3 L = primal ? V3(0.f) : state_in, throughput(1.f); float eta = 1.f; bool active = true;
    SurfaceInteraction si; bool needs_intersection = true; Interaction last_scatter_event; float last_scatter_direction_pdf = 1.f;
    int medium = -1;                                                /* :127-128 sensors inside media unsupported */
    bool valid_ray = false, specular_chain = true;
    uint32_t channel = (uint32_t) std::min(3.f * smp.next_1d(), 2.f);
    const float sgn = primal ? 1.f : -1.f;
    while (active) {
        active &= any_nonzero(throughput);
        float q = std::min(hmax(throughput) * sqr(eta), 0.99f);      /* :146 */
        bool perform_rr = depth > (uint32_t) rr_depth;
        if (active) { float u = smp.next_1d(); active &= (u < q) || !perform_rr; }
        if (perform_rr) throughput *= 1.f / q;
        if (!active) break;
        bool active_medium = medium >= 0, active_surface = !active_medium;
        MediumInteraction mei; V3 weight(1.f); bool escaped_medium = false, act_null_scatter = false, act_medium_scatter = false;
        V3 dweight(0.f);      /* d weight_c / d grid value at mei.p */
        V3 dweight_albedo(0.f); /* d weight_c / d albedo_c at mei.p (real scatters) */
        float scatter_prob = 1.f;
        if (active_medium) {
            const Medium &m = sc.media[medium];
            float u = smp.next_1d();
            mei = sample_interaction(medium, ray, u, channel);
            if (m.m.kind == LR_MEDIUM_HOMOGENEOUS && mei.is_valid()) ray.maxt = mei.t;
            if (needs_intersection) si = intersect(ray);
            needs_intersection = false;
            if (si.t < mei.t) mei.t = Inf;
            float t = std::min(mei.t, si.t) - mei.mint;
            V3 tr = exp3(mei.combined_extinction * (-t));
            V3 ffp = si.t < mei.t ? tr : tr * mei.combined_extinction;
            float tr_pdf = index_spectrum(ffp, channel);
            weight *= tr_pdf > 0.f ? tr / tr_pdf : V3(0.f);
            escaped_medium = !mei.is_valid(); active_medium = mei.is_valid();
            if (prb_handle_null) {
                if (active_medium) {
                    scatter_prob = hmean(mei.sigma_t / mei.combined_extinction);
                    act_null_scatter = smp.next_1d() >= scatter_prob;
                    act_medium_scatter = !act_null_scatter;
                    if (act_null_scatter) {
                        V3 A = weight;
                        weight *= mei.sigma_n / (1.f - scatter_prob);
                        dweight = A * (-m.m.scale / (1.f - scatter_prob));
                    }
                }
            } else act_medium_scatter = active_medium;
            if (act_medium_scatter) { depth += 1; last_scatter_event = mei; cnt.real_scatters++; }
        }
        active &= depth < max_depth;
        act_medium_scatter &= active;
        if (act_null_scatter) { ray.o = mei.p; si.t = si.t - mei.t; }
        if (act_medium_scatter) {
            const Medium &m = sc.media[mei.medium];
            V3 A = weight;
            weight *= mei.sigma_s / scatter_prob;
            V3 albedo = m.m.albedo_grid >= 0 ? sc.grids[m.m.albedo_grid].eval3(mei.p)
                                             : V3(m.m.albedo_const[0], m.m.albedo_const[1], m.m.albedo_const[2]);
            dweight = A * albedo * (m.m.scale / scatter_prob);
            dweight_albedo = A * mei.sigma_t / scatter_prob;
        }
        throughput *= weight;
        if (!primal && (active_medium || escaped_medium) && (act_null_scatter || act_medium_scatter)) {
            /* Lo = L / max(1e-8, weight); backward(dL * weight * Lo)  (:200-204) */
            double g = 0.0;
            for (int c = 0; c < 3; ++c) g += (double) (dL[c] * (L[c] / std::max(1e-8f, weight[c]))) * (double) dweight[c];
            grid_grad_add(mei, g);
            if (act_medium_scatter) {
                V3 ga;
                for (int c = 0; c < 3; ++c) ga[c] = (dL[c] * (L[c] / std::max(1e-8f, weight[c]))) * dweight_albedo[c];
                albedo_grad_add(mei, ga);
            }
        }
        Phase phase = act_medium_scatter ? phase_of(mei.medium) : Phase{LR_PHASE_ISOTROPIC, 0.f};

        active_surface |= escaped_medium;
        bool intersect_now = active_surface && needs_intersection;
        if (intersect_now) si = intersect(ray);
        if (hide_emitters && intersect_now && depth == 0u && si.is_valid() && sc.meshes[si.mesh].emitter >= 0) {
            /* hide area emitters on the first segment (prbvolpath.py:218-231: si recomputed with the ORIGINAL ray) */
            Ray r2 = si.spawn_ray(ray.d); Hit pi = skip_area_emitters(r2);
            si = sc.compute_si(ray, pi);
        }
        if (active_surface) {                                                     /* :236-252 */
            bool count_direct = depth == 0u || specular_chain;
            int emitter = sc.emitter_of(si);
            bool active_e = emitter >= 0 && !(depth == 0u && hide_emitters);
            if (active_e) {
                DirectionSample ds = em.ds_from_si(si, last_scatter_event);
                float emitter_pdf = em.pdf_emitter_direction(last_scatter_event, ds);
                V3 emitted = em.eval(emitter, si);
                V3 contrib = count_direct ? throughput * emitted : throughput * mis_weight(last_scatter_direction_pdf, emitter_pdf) * emitted;
                L += contrib * sgn;
            }
        }
        active_surface &= si.is_valid();
        /* ---- NEE (:259-291) ---- */
        {
            bool active_e_surface = false;
            if (active_surface) { BSDF b(sc, si); active_e_surface = (b.flags() & F_SMOOTH) && (depth + 1 < max_depth); }
            bool sample_emitters = act_medium_scatter ? sc.media[mei.medium].m.sample_emitters != 0 : false;
            if (act_medium_scatter) { specular_chain = false; specular_chain |= !sample_emitters; }
            bool active_e_medium = act_medium_scatter && sample_emitters;
            if (active_e_surface || active_e_medium) {
                Sampler nee_sampler = smp;
                DirectionSample ds;
                const Interaction &ref = active_e_medium ? (const Interaction &) mei : (const Interaction &) si;
                V3 emitted = sample_emitter_prb(ref, active_e_surface ? &si : nullptr, smp, medium, channel, ds, nullptr, V3(0.f));
                V3 nee_weight; float nee_pdf;
                if (active_e_surface) { BSDF b(sc, si); V3 wo = si.to_local(ds.d); nee_weight = b.eval(si.wi, wo); nee_pdf = b.pdf(si.wi, wo); }
                else { float pv = phase.eval_pdf(mei.wi, ds.d); nee_weight = V3(pv); nee_pdf = pv; }
                V3 contrib = throughput * nee_weight * mis_weight(ds.pdf, ds.delta ? 0.f : nee_pdf) * emitted;
                L += contrib * sgn;
                if (!primal) sample_emitter_prb(ref, active_e_surface ? &si : nullptr, nee_sampler, medium, channel, ds, &contrib, dL);
            }
        }
        /* ---- phase sampling (:295-314) ---- */
        valid_ray |= act_medium_scatter;
        if (act_medium_scatter) {
            float s1 = smp.next_1d(); (void) s1; V2 s2 = smp.next_2d();
            float phase_pdf; V3 wo = phase.sample(mei.wi, s2, phase_pdf);
            act_medium_scatter &= phase_pdf > 0.f;
            if (act_medium_scatter) { ray = mei.spawn_ray_m(wo); needs_intersection = true; last_scatter_direction_pdf = phase_pdf; }
        }
        /* ---- BSDF sampling (:318-349) ---- */
        if (active_surface) {
            BSDF bsdf(sc, si);
            float s1 = smp.next_1d(); V2 s2 = smp.next_2d();
            V3 bsdf_weight; BSDFSample bs = bsdf.sample(si.wi, s1, s2, bsdf_weight);
            active_surface &= bs.pdf > 0.f;
            if (active_surface) {
                throughput *= bsdf_weight; eta *= bs.eta;
                ray = si.spawn_ray(si.to_world(bs.wo));
                needs_intersection = true;
                bool non_null = !(bs.sampled_type & F_NULL);
                if (non_null) { depth += 1; last_scatter_event = si; last_scatter_direction_pdf = bs.pdf; }
                valid_ray |= non_null;
                specular_chain |= non_null && (bs.sampled_type & F_DELTA);
                specular_chain &= !(bs.sampled_type & F_SMOOTH);
                if (sc.meshes[si.mesh].is_medium_transition()) medium = sc.target_medium(si, ray.d);
            }
        }
        active &= (active_surface || active_medium);
    }
    return {L, valid_ray};
}
