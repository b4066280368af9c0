/* TEST INFRASTRUCTURE ONLY.  Included inside struct orc::Integrator.
 *
 * The fork's tissue media (src/media/liver.cpp:133-574, glissonCapsule.cpp:134-414, parenchyma.cpp:134-393) and
 * its biovolpath integrator (src/integrators/biovolpath.cpp:95-551), one lane, with the masking semantics of the
 * JIT variants (cuda_rgb / llvm_ad_rgb):
 *   - `if (dr::any_or<true>(c)) A; else if (dr::any_or<true>(d)) B; ...` runs A only (any_or<true> is `true` for a
 *     JIT array), every statement of A under its own per-lane mask.  liver.cpp uses independent `if`s, so it is the
 *     same in every variant; glissonCapsule.cpp's chain only matters where `active` is false, which it never is
 *     (:369); parenchyma.cpp's chains (:333-341, :345-356) differ from scalar_rgb: an absorbed lane keeps
 *     transmittance 1 and the hepatocyte-diameter test never runs.  Restated as the JIT variants execute it.
 *   - a symbolic dr::while_loop keeps every state write of an iteration for the lanes whose condition held when the
 *     iteration began, so statements without a per-lane mask also take effect in the iteration in which a lane
 *     dies (Russian roulette / max_depth at the top): biovolpath.cpp:222 assigns `mei` from a virtual call whose
 *     mask Dr.Jit does not pick up (the Mask is not the last argument) -- lanes without a medium get zeros
 *     (transmittance 0), lanes with one get a fresh interaction from the sampler's NEXT value, which a masked
 *     next_1d returns without consuming it -- and :304 then zeroes `result` per channel where that transmittance is 0.
 * PARITY UNPINNED: the fork's media do not compile against the vendored headers and the
 * reference ships no test or golden image for them; this file is pinned to the source text only.
 */
enum { BioAbsorber = 0, BioAttenuator = 1, BioAbsorberAndAttenuator = 2 };   /* organic_material.h:30-35 */

V3 bio_sigma_t(const Medium &m) const {
    if (m.m.kind == LR_MEDIUM_PARENCHYMA) return V3(77.2f / 255, 105.0f / 255, 149.0f / 255);   /* parenchyma.cpp:164,182 */
    return V3(m.m.sigma_t_const[0], m.m.sigma_t_const[1], m.m.sigma_t_const[2]) * m.m.scale;     /* liver.cpp:206 */
}

/* computeDistance (liver.cpp:229-481, glissonCapsule.cpp:236-305, parenchyma.cpp:193-303) */
static float bio_compute_distance(const lr_medium &m, uint32_t channel, float sample, float depth, int &bio_type) {
    uint32_t seed; memcpy(&seed, &sample, 4);
    PCG32 rng; rng.seed((uint64_t) seed);                          /* liver.cpp:235-237, default stream */
    float distance = Inf; int element = 0; bio_type = 0;
    int layer = 0;
    if (m.kind != LR_MEDIUM_PARENCHYMA) {                          /* liver.cpp:248-253: later assignments win */
        if (depth <= m.layer_limit[0]) layer = 0;
        if (depth <= m.layer_limit[1]) layer = 1;
        if (depth <= m.layer_limit[2]) layer = 2;
        if (depth <= m.layer_limit[3]) layer = 3;
        if (depth >  m.layer_limit[3]) layer = 4;
    } else layer = 4;
    if (layer < 4) {                                               /* capsule: collagen, elastin (liver.cpp:309-343) */
        for (int i = 0; i < 2; ++i) {
            float r = rng.next_f32(); if (r == 0.f) r = .5f;
            const float *s = i == 0 ? m.sigma_collagen[layer] : m.sigma_elastin[layer];
            float att = s[channel == 1u ? 1 : (channel == 2u ? 2 : 0)];
            float aux = (-(1.f / att)) * logf(r);
            if (i == 0 || aux < distance) { element = i; distance = aux; }
        }
        bio_type = BioAttenuator;
    } else if (m.kind != LR_MEDIUM_GLISSON) {                      /* parenchyma elements (liver.cpp:345-394) */
        static const int types[4] = {BioAbsorber, BioAbsorber, BioAbsorber, BioAbsorberAndAttenuator};
        for (int i = 0; i < 4; ++i) {
            float r = rng.next_f32(); if (r == 0.f) r = .5f;
            const float hep[3] = {m.sigma_hepatocity, m.sigma_hepatocity, m.sigma_hepatocity};
            const float *s = i == 0 ? m.sigma_blood : (i == 1 ? m.sigma_bile : (i == 2 ? m.sigma_lipid_water : hep));
            float att = s[(i != 3 && channel == 1u) ? 1 : ((i != 3 && channel == 2u) ? 2 : 0)];
            if (att > 0.f) {
                float aux = (-(1.f / att)) * logf(r);
                float log10v = log2f(att + 1.f) * (1.f / log2f(10.f));
                if (types[i] == BioAbsorberAndAttenuator) aux = -(log10v * logf(r));
                if (i == 0 || aux < distance) { element = i; distance = aux; }
            }
        }
        bio_type = types[element];
    } else bio_type = BioAttenuator;                               /* glissonCapsule.cpp:281-283, loop body runs masked */
    return distance;
}

/* sample_interaction with the tissue depth (liver.cpp:483-547, glissonCapsule.cpp:307-383, parenchyma.cpp:305-362);
   `maxt` is what the integrator passes as Ray3f(ray, si.t).maxt */
MediumInteraction bio_sample_interaction(int mid, const Ray &ray, float maxt, float sample, uint32_t channel, float depth, bool active_in = true) {
    const Medium &md = sc.media[mid]; const lr_medium &m = md.m;
    MediumInteraction mei; mei.wi = -ray.d; mei.medium = mid; mei.n = V3(0.f); mei.p = V3(0.f); mei.mint = 0.f;
    V3 ce = active_in ? bio_sigma_t(md) : V3(0.f);                /* get_majorant(mei, active) */
    int bt; float distance = bio_compute_distance(m, channel, sample, depth, bt);
    float dist_surf = maxt;
    bool valid_mi = active_in && distance <= maxt;
    mei.transmittance = V3(1.f);
    bool mask = distance > 0.f && distance < dist_surf;
    if (mask) { mei.t = distance; mei.p = ray_at(ray, mei.t); }
    bool active = active_in;      /* false only in the iteration in which the lane dies, see the header */
    V3 one_hot = channel == 1u ? V3(0.f, 1.f, 0.f) : (channel == 2u ? V3(0.f, 0.f, 1.f) : V3(1.f, 0.f, 0.f));
    if (m.kind == LR_MEDIUM_LIVER) {
        if (bt == BioAbsorber) active = false;
        if (bt == BioAttenuator) active = true;
        if (bt != BioAbsorber && bt != BioAttenuator && (double) distance < 0.0025) active = false;   /* Float64 r, liver.cpp:517-520 */
        if (mask && active) mei.transmittance = one_hot;
        if (mask && !active) mei.transmittance = V3(0.f);
        if (!mask) { mei.transmittance = V3(1.f); mei.t = Inf; }
    } else if (m.kind == LR_MEDIUM_GLISSON) {
        active = true;                                               /* glissonCapsule.cpp:369 */
        if (mask && active) mei.transmittance = one_hot;             /* first branch of the chain only; t is Inf from zeros<> */
    } else {
        if (bt == BioAbsorber) active = false;                       /* first branch of the chain only (parenchyma.cpp:333) */
        if (mask && active) mei.transmittance = one_hot;             /* first branch only (:345-350): absorbed lanes keep 1 */
    }
    cnt.collisions++;
    mei.sigma_t = valid_mi ? bio_sigma_t(md) : V3(0.f);
    mei.sigma_s = m.kind == LR_MEDIUM_PARENCHYMA ? (valid_mi ? V3(74.0f / 255, 88.0f / 255, 101.0f / 255) : V3(0.f))
                                                 : mei.sigma_t * V3(m.albedo_const[0], m.albedo_const[1], m.albedo_const[2]);
    mei.sigma_n = V3(0.f);
    mei.combined_extinction = ce;
    return mei;
}

/* BioVolumetricPathIntegrator::sample_emitter (biovolpath.cpp:399-531) */
V3 sample_emitter_bio(const Interaction &ref, const SurfaceInteraction *ref_si, Sampler &smp, int medium,
                      uint32_t channel, float tissue_depth, DirectionSample &ds_out) {
    V3 transmittance(1.f), emitter_val;
    DirectionSample ds = em.sample_emitter_direction(ref, smp.next_2d(), false, emitter_val);
    if (ds.pdf == 0.f) emitter_val = V3(0.f);
    bool active = ds.pdf != 0.f;
    ds_out = ds;
    if (!active) return emitter_val;

    Ray ray = ref.spawn_ray_to(ds.p);
    float max_dist = ray.maxt;
    if (ref_si && sc.meshes[ref_si->mesh].is_medium_transition()) medium = sc.target_medium(*ref_si, ray.d);

    float total_dist = 0.f; SurfaceInteraction si;
    bool needs_intersection = true;
    while (active) {
        float remaining_dist = max_dist - total_dist;
        ray.maxt = remaining_dist;
        active &= remaining_dist > 0.f;
        if (!active) break;

        bool escaped_medium = false, active_medium = medium >= 0, active_surface = !active_medium;
        if (active_medium) {
            const Medium &m = sc.media[medium];
            MediumInteraction mei = bio_sample_interaction(medium, ray, si.t, smp.next_1d(), channel, tissue_depth);   /* :472-474 */
            if (mei.is_valid()) ray.maxt = std::min(mei.t, remaining_dist);     /* every tissue medium is_homogeneous */
            if (needs_intersection) si = intersect(ray);
            if (si.t < mei.t) mei.t = Inf;
            needs_intersection = false;
            bool is_spectral = m.m.has_spectral_extinction != 0;
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
            if (active_medium) {
                total_dist += mei.t;
                ray.o = mei.p; si.t = si.t - mei.t;
                transmittance *= mei.transmittance;                              /* :501-504, both branches */
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

/* biovolpath.cpp:304-305 for one lane: `mei` is the lane's loop-state MediumInteraction */
static void bio_apply_transmittance(const MediumInteraction &mei, int medium, V3 &result, V3 &throughput) {
    if (mei.transmittance.x == 0.f) result.x = 0.f;
    if (mei.transmittance.y == 0.f) result.y = 0.f;
    if (mei.transmittance.z == 0.f) result.z = 0.f;
    if (medium >= 0) throughput *= mei.transmittance;
}

/* BioVolumetricPathIntegrator::sample (biovolpath.cpp:95-396) */
LaneResult sample_biovolpath(Sampler &smp, Ray ray, int initial_medium) {
    bool valid_ray = !hide_emitters && sc.env >= 0;
    float eta = 1.f; V3 throughput(1.f), result(0.f);
    int medium = initial_medium;
    MediumInteraction mei; mei.transmittance = V3(0.f); mei.p = V3(0.f);          /* dr::zeros<> (interaction.h:604-614) */
    bool specular_chain = !hide_emitters; uint32_t depth = 0;
    uint32_t channel = (uint32_t) std::min(smp.next_1d() * 3.f, 2.f);
    SurfaceInteraction si;
    bool needs_intersection = true;
    Interaction last_scatter_event;
    float last_scatter_direction_pdf = 1.f, tissue_depth = 0.f;
    bool active = true;
    while (active) {
        active &= any_nonzero(throughput);                                          /* :194-200 */
        float q = std::min(hmax(throughput) * sqr(eta), .95f);
        bool perform_rr = depth > (uint32_t) rr_depth;
        if (active) { float u = smp.next_1d(); active &= (u < q) || !perform_rr; }
        if (perform_rr) throughput *= 1.f / q;
        active &= depth < max_depth;
        if (!active) {                                          /* the rest of the body is masked; :222 and :304 are not */
            if (medium >= 0) { Sampler peek = smp; mei = bio_sample_interaction(medium, ray, si.t, peek.next_1d(), channel, tissue_depth, false); }
            else { mei = MediumInteraction(); mei.transmittance = V3(0.f); }
            bio_apply_transmittance(mei, medium, result, throughput); break;
        }

        bool active_medium = medium >= 0, active_surface = !active_medium;
        bool act_null_scatter = false, act_medium_scatter = false, escaped_medium = false;
        bool is_spectral = active_medium && sc.media[medium].m.has_spectral_extinction != 0;

        if (active_medium) {
            mei = bio_sample_interaction(medium, ray, si.t, smp.next_1d(), channel, tissue_depth);   /* :222 */
            if (mei.is_valid()) ray.maxt = mei.t;
            if (needs_intersection) si = intersect(ray);
            needs_intersection = false;
            if (si.t < mei.t) mei.t = Inf;
            if (is_spectral) {                                                      /* :231-235, medium.cpp:95-104 */
                float t = std::min(mei.t, si.t) - mei.mint;
                V3 tr = exp3(mei.combined_extinction * (-t));
                V3 pdf = si.t < mei.t ? tr : tr * mei.combined_extinction;
                float tr_pdf = index_spectrum(pdf, channel);
                throughput *= tr_pdf > 0.f ? tr / tr_pdf : V3(0.f);
            }
            escaped_medium = !mei.is_valid();
            active_medium = mei.is_valid();
            if (active_medium) {
                bool null_scatter = smp.next_1d() >= index_spectrum(mei.sigma_t, channel) / index_spectrum(mei.combined_extinction, channel);   /* :241 */
                act_null_scatter = null_scatter; act_medium_scatter = !null_scatter;
                if (is_spectral && act_null_scatter)
                    throughput *= mei.sigma_n * index_spectrum(mei.combined_extinction, channel) / index_spectrum(mei.sigma_n, channel);
                if (act_medium_scatter) { depth += 1; last_scatter_event = mei; cnt.real_scatters++; }
            }
        } else { mei = MediumInteraction(); mei.transmittance = V3(0.f); }          /* :222, lanes with medium == nullptr */
        active &= depth < max_depth;                                                /* :256 */
        act_medium_scatter &= active;
        if (act_null_scatter) { ray.o = mei.p; si.t = si.t - mei.t; }

        if (act_medium_scatter) {
            throughput *= mei.transmittance;                                        /* :266-275 (both branches alike) */
            tissue_depth += fabsf((-ray.d).z * mei.t);
            Phase phase = phase_of(mei.medium);
            float s1 = smp.next_1d(); (void) s1; V2 s2 = smp.next_2d();              /* :283-286 */
            float phase_pdf; V3 wo = phase.sample(mei.wi, s2, phase_pdf);
            if (phase_pdf > 0.f) {
                ray = mei.spawn_ray_m(wo);
                needs_intersection = true;
                last_scatter_direction_pdf = phase_pdf;
            }
        }

        /* ---- surfaces (:298-388) ---- */
        active_surface |= escaped_medium;
        bool intersect_now = active_surface && needs_intersection;
        bio_apply_transmittance(mei, medium, result, throughput);                   /* :303-306 */
        if (intersect_now) si = intersect(ray);
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
            if (active_e) {                                                          /* :342-356 */
                DirectionSample ds;
                V3 emitted = sample_emitter_bio(si, &si, smp, medium, channel, tissue_depth, ds);
                V3 wo = si.to_local(ds.d);
                V3 bsdf_val = bsdf.eval(si.wi, wo); float bsdf_pdf = bsdf.pdf(si.wi, wo);
                result += throughput * bsdf_val * mis_weight(ds.pdf, ds.delta ? 0.f : bsdf_pdf) * emitted;
            }
            float s1 = smp.next_1d(); V2 s2 = smp.next_2d();
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
