// lr_biovolpath.cuh -- the fork's tissue media and its biovolpath integrator (device code, one lane per thread).
// Media: src/media/liver.cpp:133-574, glissonCapsule.cpp:134-414, parenchyma.cpp:134-393 (element types:
// organic_material.h:30-35).  Integrator: src/integrators/biovolpath.cpp:95-551.
// Semantics are those of the reference's JIT variants (cuda_rgb / llvm_ad_rgb), including the statements that run without
// a per-lane mask -- see the notes at bio_sample_interaction and above BioPath.  Every tissue medium is
// `m_is_homogeneous` with sigma_n = 0: a free path is not an exponential in sigma_t but the minimum of one exponential per
// tissue element, drawn from an inner PCG32 seeded with the bit pattern of the integrator's sample.
#pragma once
#include "lr_integrators.cuh"

namespace lr {

enum { BIO_ABSORBER = 0, BIO_ATTENUATOR = 1, BIO_ABSORBER_AND_ATTENUATOR = 2 };

struct BioInteraction { float t; V3 p, wi, transmittance, combined_extinction; float sigma_t_ch; };
LR_DEV bool bio_valid(const BioInteraction &m) { return m.t != LR_INF; }
LR_DEV BioInteraction bio_zeros() {          // dr::zeros<MediumInteraction3f>() (interaction.h:604-614): t = inf, everything else 0
    BioInteraction m; m.t = LR_INF; m.p = m.wi = m.transmittance = m.combined_extinction = mk3(0.f); m.sigma_t_ch = 0.f; return m;
}

LR_DEV V3 bio_sigma_t(const lr_medium &m) {
    if (m.kind == LR_MEDIUM_PARENCHYMA) return mk3(77.2f / 255, 105.0f / 255, 149.0f / 255);     // parenchyma.cpp:164,182
    return mk3(m.sigma_t_const[0], m.sigma_t_const[1], m.sigma_t_const[2]) * m.scale;             // liver.cpp:206
}

struct BioRng {                              // dr::PCG32<Float>::seed(initstate = seed, default stream) + next_float32
    uint64_t state;
    static constexpr uint64_t kInc = (0xda3e39cb94b95bdbULL << 1) | 1ull;
    LR_DEV uint32_t next_u32() {
        uint64_t old = state; state = old * 0x5851f42d4c957f2dULL + kInc;
        uint32_t xs = (uint32_t) (((old >> 18u) ^ old) >> 27u), rot = (uint32_t) (old >> 59u);
        return __funnelshift_r(xs, xs, rot);
    }
    LR_DEV void seed(uint32_t s) { state = 0; next_u32(); state += (uint64_t) s; next_u32(); }
    LR_DEV float next_f32() { return __uint_as_float((next_u32() >> 9) | 0x3f800000u) - 1.f; }
};

// computeDistance (liver.cpp:229-481, glissonCapsule.cpp:236-305, parenchyma.cpp:193-303)
LR_DEV float bio_compute_distance(const lr_medium &m, uint32_t channel, float sample, float depth, int &bio_type) {
    BioRng rng; rng.seed(__float_as_uint(sample));
    float distance = LR_INF; int element = 0; bio_type = 0;
    int layer = 4;
    if (m.kind != LR_MEDIUM_PARENCHYMA) {                         // liver.cpp:248-253: later assignments win
        layer = 0;
        if (depth <= m.layer_limit[0]) layer = 0;
        if (depth <= m.layer_limit[1]) layer = 1;
        if (depth <= m.layer_limit[2]) layer = 2;
        if (depth <= m.layer_limit[3]) layer = 3;
        if (depth >  m.layer_limit[3]) layer = 4;
    }
    const int c = channel == 1u ? 1 : (channel == 2u ? 2 : 0);
    if (layer < 4) {                                              // capsule: collagen, elastin (liver.cpp:309-343)
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            float r = rng.next_f32(); if (r == 0.f) r = .5f;
            float att = i == 0 ? m.sigma_collagen[layer][c] : m.sigma_elastin[layer][c];
            float aux = __fmul_rn(-(1.f / att), logf(r));
            if (i == 0 || aux < distance) { element = i; distance = aux; }
        }
        bio_type = BIO_ATTENUATOR;
    } else if (m.kind != LR_MEDIUM_GLISSON) {                     // parenchyma elements (liver.cpp:345-394)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float r = rng.next_f32(); if (r == 0.f) r = .5f;
            float att = i == 0 ? m.sigma_blood[c] : (i == 1 ? m.sigma_bile[c] : (i == 2 ? m.sigma_lipid_water[c] : m.sigma_hepatocity));
            if (att > 0.f) {
                float lr_ = logf(r);
                float aux = __fmul_rn(-(1.f / att), lr_);
                if (i == 3) aux = -__fmul_rn(__fmul_rn(log2f(att + 1.f), 1.f / log2f(10.f)), lr_);   // EAbsorberAndAttenuator
                if (i == 0 || aux < distance) { element = i; distance = aux; }
            }
        }
        bio_type = element == 3 ? BIO_ABSORBER_AND_ATTENUATOR : BIO_ABSORBER;
    } else bio_type = BIO_ATTENUATOR;                             // glissonCapsule.cpp:281-283 (the loop body runs masked)
    return distance;
}

// sample_interaction with the tissue depth (liver.cpp:483-547, glissonCapsule.cpp:307-383, parenchyma.cpp:305-362).
// `maxt` is what the integrator passes as Ray3f(ray, si.t).maxt.  `active_in` is the Mask argument: false only in the
// iteration in which a lane dies (note (2) above BioPath).  JIT semantics of the `if (any_or<true>(..)) .. else if ..`
// chains in glissonCapsule.cpp / parenchyma.cpp: only the first branch executes (under its own mask).
LR_DEV BioInteraction bio_sample_interaction(const lr_medium &m, const Ray &ray, float maxt, float sample, uint32_t channel, float depth,
                                             bool active_in, uint32_t &n_collisions) {
    BioInteraction mei = bio_zeros(); mei.wi = -ray.d;
    const V3 st = bio_sigma_t(m);
    int bt; const float distance = bio_compute_distance(m, channel, sample, depth, bt);
    const bool valid_mi = active_in && distance <= maxt;
    const bool mask = distance > 0.f && distance < maxt;
    mei.transmittance = mk3(1.f);
    if (mask) { mei.t = distance; mei.p = ray_at(ray, distance); }
    bool active = active_in;
    const V3 one_hot = mk3(channel == 0u ? 1.f : 0.f, channel == 1u ? 1.f : 0.f, channel == 2u ? 1.f : 0.f);
    if (m.kind == LR_MEDIUM_LIVER) {
        if (bt == BIO_ABSORBER) active = false;
        if (bt == BIO_ATTENUATOR) active = true;
        if (bt == BIO_ABSORBER_AND_ATTENUATOR && (double) distance < 0.0025) active = false;       // Float64 r (liver.cpp:517-520)
        if (mask) mei.transmittance = active ? one_hot : mk3(0.f);
    } else if (m.kind == LR_MEDIUM_GLISSON) {
        if (mask) mei.transmittance = one_hot;                                                     // active = true (:369)
    } else {
        if (bt == BIO_ABSORBER) active = false;                                                    // parenchyma.cpp:333
        if (mask && active) mei.transmittance = one_hot;                                           // :345-350, absorbed lanes keep 1
    }
    n_collisions++;
    mei.sigma_t_ch = valid_mi ? idx3(st, channel) : 0.f;
    mei.combined_extinction = active_in ? st : mk3(0.f);
    return mei;
}

// The walk of BioVolumetricPathIntegrator::sample_emitter (biovolpath.cpp:444-528) along the shadow ray `ray` (maxt = max_dist).
// `first_hit`: closest hit along the WHOLE shadow segment when a trace stage has already found it (wavefront), else nullptr.  The
// reference shortens the ray to the sampled free path before it intersects (:476-480), so a hit beyond that distance is not seen
// -- and, as no later iteration intersects again, never will be: the walk then runs through the boundary to the emitter.  With a
// precomputed hit that is one comparison.
LR_DEV V3 bio_nee_walk(const DScene &sc, const DTri *tris_gid, Ray ray, float ds_dist, int medium, uint32_t channel, float tissue_depth,
                       const Hit *first_hit, Sampler &smp, Counters &cnt) {
    V3 transmittance = mk3(1.f);
    const float max_dist = ray.maxt;
    float total_dist = 0.f; SurfaceInteraction si = invalid_si(); bool needs_intersection = true, active = true;
    auto isect = [&](const Ray &r) {
        if (first_hit) { Hit h = *first_hit; first_hit = nullptr; if (!(h.t <= r.maxt)) { h.t = LR_INF; h.prim = 0xffffffffu; } return compute_si(sc, r, h, tris_gid); }
        return intersect(sc, tris_gid, r, cnt);
    };
    while (active) {
        const float remaining_dist = max_dist - total_dist;
        ray.maxt = remaining_dist;
        active &= remaining_dist > 0.f;
        if (!active) break;
        bool escaped_medium = false, active_medium = medium >= 0, active_surface = !active_medium;
        if (active_medium) {
            const lr_medium &m = sc.media[medium].m;
            BioInteraction mei = bio_sample_interaction(m, ray, si.t, smp.next_1d(), channel, tissue_depth, true, cnt.collisions);
            if (bio_valid(mei)) ray.maxt = fminf(mei.t, remaining_dist);
            if (needs_intersection) si = isect(ray);
            if (si.t < mei.t) mei.t = LR_INF;
            needs_intersection = false;
            if (m.has_spectral_extinction) {
                float t = fminf(remaining_dist, fminf(mei.t, si.t));
                V3 tr = exp3(mei.combined_extinction * (-t));
                V3 ffp = (si.t < mei.t || mei.t > remaining_dist) ? tr : tr * mei.combined_extinction;
                float tr_pdf = idx3(ffp, channel);
                transmittance *= tr_pdf > 0.f ? tr / tr_pdf : mk3(0.f);
            }
            if (mei.t > remaining_dist && bio_valid(mei)) total_dist = ds_dist;
            if (mei.t > remaining_dist) mei.t = LR_INF;
            escaped_medium = !bio_valid(mei); active_medium = bio_valid(mei);
            if (active_medium) {
                total_dist += mei.t; ray.o = mei.p; si.t = si.t - mei.t;
                transmittance *= mei.transmittance;                                    // :501-504
            }
        }
        if (active_surface && needs_intersection) { si = isect(ray); needs_intersection = false; }
        active_surface |= escaped_medium;
        if (active_surface) total_dist += si.t;
        active_surface &= si_valid(si) && !active_medium;
        if (active_surface) {
            BSDF bsdf(sc, si);
            transmittance *= bsdf.eval_null_transmission();
            Ray r2 = spawn_ray(si.p, si.n, ray.d); ray.o = r2.o;
        }
        ray.maxt = remaining_dist;
        needs_intersection |= active_surface;
        active &= (active_medium || active_surface) && any_nonzero(transmittance);
        if (active_surface && mesh_is_transition(sc, si.mesh)) medium = target_medium(sc, si, ray.d);
    }
    return transmittance;
}

// BioVolumetricPathIntegrator::sample_emitter (biovolpath.cpp:399-531)
LR_DEV V3 sample_emitter_bio(const DScene &sc, const DTri *tris_gid, const SurfaceInteraction &ref, Sampler &smp, int medium,
                             uint32_t channel, float tissue_depth, DirectionSample &ds, Counters &cnt) {
    V3 emitter_val;
    ds = sample_emitter_direction(sc, ref.p, smp.next_2d(), emitter_val);
    if (ds.pdf == 0.f) return mk3(0.f);
    const Ray ray = spawn_ray_to(ref.p, ref.n, ds.p);
    if (mesh_is_transition(sc, ref.mesh)) medium = target_medium(sc, ref, ray.d);
    return bio_nee_walk(sc, tris_gid, ray, ds.dist, medium, channel, tissue_depth, nullptr, smp, cnt) * emitter_val;
}

// biovolpath.cpp:304-305 for one lane: `mei` is the lane's loop-state MediumInteraction
LR_DEV void bio_apply_transmittance(const BioInteraction &mei, int medium, V3 &result, V3 &throughput) {
    if (mei.transmittance.x == 0.f) result.x = 0.f;
    if (mei.transmittance.y == 0.f) result.y = 0.f;
    if (mei.transmittance.z == 0.f) result.z = 0.f;
    if (medium >= 0) throughput *= mei.transmittance;
}

// BioVolumetricPathIntegrator::sample (biovolpath.cpp:95-396).  Two things a symbolic dr::while_loop does that a scalar
// reading of the source would miss -- it keeps EVERY state write of an iteration for the lanes whose loop condition held
// when the iteration began:
//  (1) :222 assigns `mei` from a virtual call whose mask Dr.Jit does not pick up (the Mask is not the last argument), so
//      lanes without a medium get zeros (transmittance 0) every iteration;
//  (2) in the iteration in which a lane dies at the top (Russian roulette / max_depth), lanes inside a medium still get a
//      fresh interaction -- from the sampler's NEXT value, which a masked next_1d returns without consuming it, and with
//      active = false -- and :304 zeroes `result` per channel where that interaction's transmittance is 0.
// The loop is written as a resumable path: init() + one step() per iteration of the reference's loop -- the unit the production wavefront
// (lr_wavefront_bio.cuh) cuts at; this per-lane form is the parity probe.  Handing a finished lane a new path between steps was tried
// first and bought nothing (0.98x / 0.88x on the two liver scenes): the idle lanes of k_render<biovolpath> -- 8 of 32 threads active per
// instruction, profiles/r2_kernel_metrics_bio_per_lane.txt -- sit inside a step (BVH walks and shadow walks of different lengths).
struct BioPath {
    Sampler smp; Ray ray; V3 throughput, result, last_scatter_p;
    float eta, last_scatter_direction_pdf, tissue_depth, si_t;            // si_t: the previous intersection's distance (what :222 passes as maxt)
    int medium; uint32_t depth, channel; bool valid_ray, specular_chain;

    LR_DEV void init(const DScene &sc, const Sampler &s, const Ray &r, int initial_medium) {
        smp = s; ray = r; valid_ray = !sc.hide_emitters && sc.env >= 0;
        eta = 1.f; throughput = mk3(1.f); result = mk3(0.f); medium = initial_medium;
        specular_chain = !sc.hide_emitters; depth = 0;
        channel = (uint32_t) fminf(smp.next_1d() * 3.f, 2.f);
        si_t = LR_INF; last_scatter_p = mk3(0.f); last_scatter_direction_pdf = 1.f; tissue_depth = 0.f;
    }
    LR_DEV LaneResult finish() const { LaneResult out; out.L = result; out.valid = valid_ray; return out; }

    // one iteration of the loop; false: the lane's loop has ended.  `needs_intersection` is true at the top of every iteration
    // (every event spawns a new ray, and a null collision -- the only thing that keeps a hit -- cannot happen: sigma_n = 0).
    LR_DEV bool step(const DScene &sc, const DTri *tris_gid, Counters &cnt) {
        bool active = any_nonzero(throughput);
        float q = fminf(hmax(throughput) * sqr(eta), .95f);
        bool perform_rr = depth > (uint32_t) sc.rr_depth;
        if (active) { float u = smp.next_1d(); active &= (u < q) || !perform_rr; }
        if (perform_rr) throughput *= 1.f / q;
        active &= depth < sc.max_depth;
        BioInteraction mei;
        if (!active) {                                                                 // note (2)
            if (medium >= 0) { Sampler peek = smp; mei = bio_sample_interaction(sc.media[medium].m, ray, si_t, peek.next_1d(), channel, tissue_depth, false, cnt.collisions); }
            else mei = bio_zeros();
            bio_apply_transmittance(mei, medium, result, throughput);
            return false;
        }
        SurfaceInteraction si = invalid_si(); si.t = si_t; bool needs_intersection = true;
        bool active_medium = medium >= 0, active_surface = !active_medium;
        bool act_medium_scatter = false, escaped_medium = false, is_spectral = false;
        if (active_medium) {
            const lr_medium &m = sc.media[medium].m;
            is_spectral = m.has_spectral_extinction != 0;
            mei = bio_sample_interaction(m, ray, si.t, smp.next_1d(), channel, tissue_depth, true, cnt.collisions);   // :222
            if (bio_valid(mei)) ray.maxt = mei.t;
            si = intersect(sc, tris_gid, ray, cnt);
            needs_intersection = false;
            if (si.t < mei.t) mei.t = LR_INF;
            if (is_spectral) {                                                         // :231-235, medium.cpp:95-104
                float t = fminf(mei.t, si.t);
                V3 tr = exp3(mei.combined_extinction * (-t));
                V3 pdf = si.t < mei.t ? tr : tr * mei.combined_extinction;
                float tr_pdf = idx3(pdf, channel);
                throughput *= tr_pdf > 0.f ? tr / tr_pdf : mk3(0.f);
            }
            escaped_medium = !bio_valid(mei); active_medium = bio_valid(mei);
            if (active_medium) {
                // :241 null_scatter = u >= sigma_t[ch] / combined_extinction[ch]: the two are the same number whenever the interaction is valid
                // (valid_mi follows from 0 < distance < maxt), so the quotient is 1 -- or NaN for a zero / infinite sigma_t -- and u < 1: never true.
                // The draw is still made.
                (void) smp.next_1d();
                act_medium_scatter = true;
                if (act_medium_scatter) { depth += 1; last_scatter_p = mei.p; cnt.real_scatters++; }
            }
        } else mei = bio_zeros();                                                      // note (1)
        active &= depth < sc.max_depth;
        act_medium_scatter &= active;

        if (act_medium_scatter) {
            throughput *= mei.transmittance;                                           // :266-275
            tissue_depth += fabsf(__fmul_rn(-ray.d.z, mei.t));
            const DMedium &m = sc.media[medium]; Phase phase; phase.kind = m.m.phase; phase.g = m.m.phase_g;
            (void) smp.next_1d(); V2 s2 = smp.next_2d();
            float phase_pdf; V3 wo = phase.sample(mei.wi, s2, phase_pdf);
            if (phase_pdf > 0.f) {
                ray.o = mei.p; ray.d = wo; ray.maxt = LR_LARGEST;
                needs_intersection = true; last_scatter_direction_pdf = phase_pdf;
            }
        }

        active_surface |= escaped_medium;
        const bool intersect_now = active_surface && needs_intersection;
        bio_apply_transmittance(mei, medium, result, throughput);                      // :303-306
        if (intersect_now) si = intersect(sc, tris_gid, ray, cnt);
        if (active_surface) {
            bool count_direct = depth == 0u || specular_chain;
            int emitter = emitter_of(sc, si);
            if (emitter >= 0 && !(depth == 0u && sc.hide_emitters)) {
                float emitter_pdf = 1.f;
                if (!count_direct) { DirectionSample ds = ds_from_si(sc, si, last_scatter_p); emitter_pdf = pdf_emitter_direction(sc, ds); }
                V3 emitted = eval_emitter_at(sc, emitter, si);
                result += count_direct ? throughput * emitted : throughput * mis_weight(last_scatter_direction_pdf, emitter_pdf) * emitted;
            }
        }
        active_surface &= si_valid(si);
        if (active_surface) {
            BSDF bsdf(sc, si);
            if ((bsdf.flags() & F_SMOOTH) && (depth + 1 < sc.max_depth)) {
                DirectionSample ds;
                V3 emitted = sample_emitter_bio(sc, tris_gid, si, smp, medium, channel, tissue_depth, ds, cnt);
                V3 wo = si.sh.to_local(ds.d);
                float bsdf_pdf; V3 bsdf_val = bsdf.eval_pdf(si.wi, wo, bsdf_pdf);
                result += throughput * bsdf_val * mis_weight(ds.pdf, ds.delta ? 0.f : bsdf_pdf) * emitted;
            }
            float s1 = smp.next_1d(); V2 s2 = smp.next_2d();
            V3 bsdf_val; BSDFSample bs = bsdf.sample(si.wi, s1, s2, bsdf_val);
            throughput *= bsdf_val; eta *= bs.eta;
            ray = spawn_ray(si.p, si.n, si.sh.to_world(bs.wo));
            bool non_null = !(bs.sampled_type & F_NULL);
            if (non_null) { depth += 1; last_scatter_p = si.p; last_scatter_direction_pdf = bs.pdf; }
            valid_ray |= non_null;
            specular_chain |= non_null && (bs.sampled_type & F_DELTA);
            specular_chain &= !(bs.sampled_type & F_SMOOTH);
            if (mesh_is_transition(sc, si.mesh)) medium = target_medium(sc, si, ray.d);
        }
        si_t = si.t;
        active &= (active_surface || active_medium);
        return active;
    }
};

LR_DEV LaneResult sample_biovolpath(const DScene &sc, const DTri *tris_gid, Sampler &smp, Ray ray, int initial_medium, Counters &cnt) {
    BioPath p; p.init(sc, smp, ray, initial_medium);
    while (p.step(sc, tris_gid, cnt)) { }
    smp = p.smp;
    return p.finish();
}

} // namespace lr
