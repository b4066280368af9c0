// lr_integrators.cuh -- per-lane path construction (device code), reference draw order.
// path (src/integrators/path.cpp:95-349), volpath (src/integrators/volpath.cpp:93-554).
// One GPU thread owns one lane (= one camera path, one PCG32 stream keyed by the GLOBAL lane
// index) and follows the masked-execution semantics of the reference's JIT variants: the stream
// only advances where the reference draws with that lane's mask set.
#pragma once
#include "lr_scene.cuh"
#include "lr_bsdf.cuh"
#include "lr_emitter.cuh"
#include "lr_medium.cuh"
#include "lr_film.cuh"

namespace lr {

struct Counters { uint32_t rays, collisions, real_scatters; };
struct LaneResult { V3 L; bool valid; };

// PerspectiveCamera::sample_ray_differential (src/sensors/perspective.cpp:239-279)
LR_DEV Ray camera_ray(const DScene &sc, V2 pos_sample) {
    V3 near_p = xform_point_proj(sc.sample_to_camera, mk3(pos_sample.x, pos_sample.y, 0.f));
    V3 d = normalize(near_p);
    Ray r; r.o = mk3(sc.cam_to_world.m[3], sc.cam_to_world.m[7], sc.cam_to_world.m[11]);
    r.d = xform_vector(sc.cam_to_world, d);
    float inv_z = 1.f / d.z, near_t = sc.sensor.near_clip * inv_z, far_t = sc.sensor.far_clip * inv_z;
    r.o = r.o + r.d * near_t; r.maxt = far_t - near_t;
    return r;
}

LR_DEV SurfaceInteraction intersect(const DScene &sc, const DTri *tris_gid, const Ray &ray, Counters &cnt) {
    cnt.rays++;
    Hit h = bvh_trace<false, false>(sc, ray);
    return compute_si(sc, ray, h, tris_gid);
}
LR_DEV int hit_emitter(const DScene &sc, const DTri *tris_gid, const Hit &h) {
    return hit_valid(h) ? sc.meshes[__float_as_uint(tris_gid[h.prim].v1.w)].emitter : -1;
}
// Integrator::skip_area_emitters (src/render/integrator.cpp:97-123): first hit along `ray` that is not an
// area emitter; `ray` ends up as the segment that hit was found with.
LR_DEV Hit skip_area_emitters(const DScene &sc, const DTri *tris_gid, Ray &ray, Counters &cnt) {
    for (;;) {
        cnt.rays++; Hit pi = bvh_trace<false, false>(sc, ray);
        if (hit_emitter(sc, tris_gid, pi) < 0) return pi;
        SurfaceInteraction si = compute_si(sc, ray, pi, tris_gid);
        ray = spawn_ray(si.p, si.n, ray.d);
    }
}
LR_DEV SurfaceInteraction invalid_si() {
    SurfaceInteraction si; si.t = LR_INF; si.p = si.n = si.wi = mk3(0.f); si.sh.s = si.sh.t = si.sh.n = mk3(0.f); si.uv = mk2(0.f, 0.f); si.mesh = -1; si.prim = 0; return si;
}
LR_DEV bool mesh_is_transition(const DScene &sc, int mesh) { return (sc.meshes[mesh].flags & 1u) != 0; }
LR_DEV int target_medium(const DScene &sc, const SurfaceInteraction &si, V3 d) {       // interaction.h:322-345
    const DMesh &m = sc.meshes[si.mesh]; return dot(d, si.n) > 0.f ? m.exterior : m.interior;
}
LR_DEV int emitter_of(const DScene &sc, const SurfaceInteraction &si) { return si_valid(si) ? sc.meshes[si.mesh].emitter : sc.env; }
LR_DEV V3 eval_emitter_at(const DScene &sc, int emitter, const SurfaceInteraction &si) {
    // escaped rays carry -ray.d (world) in si.wi; area lights need cos_theta(si.wi) in the local frame
    return emitter_eval(sc, emitter, si.wi.z, -si.wi);
}

// ===================== path =====================
LR_DEV LaneResult sample_path(const DScene &sc, const DTri *tris_gid, Sampler &smp, Ray ray, Counters &cnt) {
    LaneResult out; out.L = mk3(0.f); out.valid = false;
    if (sc.max_depth == 0) return out;
    V3 throughput = mk3(1.f), result = mk3(0.f); float eta = 1.f; uint32_t depth = 0;
    bool valid_ray = !sc.hide_emitters && sc.env >= 0;
    V3 prev_p = mk3(0.f); float prev_bsdf_pdf = 1.f; bool prev_bsdf_delta = true;
    cnt.rays++; Hit pi = bvh_trace<false, false>(sc, ray);
    if (sc.hide_emitters && hit_emitter(sc, tris_gid, pi) >= 0) {         // path.cpp:178-191 (ls.ray itself is kept)
        SurfaceInteraction s0 = compute_si(sc, ray, pi, tris_gid);
        Ray r2 = spawn_ray(s0.p, s0.n, ray.d); pi = skip_area_emitters(sc, tris_gid, r2, cnt);
    }
    bool active = true;
    while (active) {
        SurfaceInteraction si = compute_si(sc, ray, pi, tris_gid);
        int emitter = emitter_of(sc, si);
        if (emitter >= 0) {
            DirectionSample ds = ds_from_si(sc, si, prev_p);
            float em_pdf = prev_bsdf_delta ? 0.f : pdf_emitter_direction(sc, ds);
            float mis_bsdf = mis_weight(prev_bsdf_pdf, em_pdf);
            V3 ev = prev_bsdf_pdf > 0.f ? eval_emitter_at(sc, emitter, si) : mk3(0.f);
            result = fma3(throughput, ev * mis_bsdf, result);
        }
        bool active_next = (depth + 1 < sc.max_depth) && si_valid(si);
        if (!si_valid(si)) break;
        BSDF bsdf(sc, si);
        bool active_em = active_next && (bsdf.flags() & F_SMOOTH);
        DirectionSample ds; ds.pdf = 0.f; ds.delta = false; ds.d = mk3(0.f); V3 em_weight = mk3(0.f), wo = mk3(0.f);
        V2 u_em = smp.next_2d();
        if (active_em) {
            ds = sample_emitter_direction(sc, si.p, u_em, em_weight);
            if (ds.pdf != 0.f) {                                  // test_visibility = true (scene.cpp:356-360)
                cnt.rays++;
                Ray sr = spawn_ray_to(si.p, si.n, ds.p);
                if (hit_valid(bvh_trace<true, false>(sc, sr))) { em_weight = mk3(0.f); ds.pdf = 0.f; }
            }
            active_em &= ds.pdf != 0.f;
            wo = si.sh.to_local(ds.d);
        }
        float s1 = smp.next_1d(); V2 s2 = smp.next_2d();
        float bsdf_pdf; V3 bsdf_val = bsdf.eval_pdf(si.wi, wo, bsdf_pdf);
        V3 bsdf_weight; BSDFSample bs = bsdf.sample(si.wi, s1, s2, bsdf_weight);
        if (active_em) {
            float mis_em = ds.delta ? 1.f : mis_weight(ds.pdf, bsdf_pdf);
            result = fma3(throughput, bsdf_val * em_weight * mis_em, result);
        }
        ray = spawn_ray(si.p, si.n, si.sh.to_world(bs.wo));
        throughput *= bsdf_weight; eta *= bs.eta;
        valid_ray |= !(bs.sampled_type & F_NULL);
        prev_p = si.p; prev_bsdf_pdf = bs.pdf; prev_bsdf_delta = (bs.sampled_type & F_DELTA) != 0;
        depth += 1;
        float tmax = hmax(throughput), rr_prob = fminf(tmax * sqr(eta), .95f);
        bool rr_active = (int) depth >= sc.rr_depth, rr_continue = smp.next_1d() < rr_prob;
        if (rr_active) throughput *= 1.f / rr_prob;
        active = active_next && (!rr_active || rr_continue) && (tmax != 0.f);
        if (active) { cnt.rays++; pi = bvh_trace<false, false>(sc, ray); }
    }
    out.L = valid_ray ? result : mk3(0.f); out.valid = valid_ray;
    return out;
}

// ===================== volpath: NEE with transmittance (volpath.cpp:400-554) =====================
LR_DEV V3 sample_emitter_vol(const DScene &sc, const DTri *tris_gid, V3 ref_p, V3 ref_n, const SurfaceInteraction *ref_si,
                             Sampler &smp, int medium, uint32_t channel, DirectionSample &ds, Counters &cnt) {
    V3 transmittance = mk3(1.f), emitter_val;
    ds = sample_emitter_direction(sc, ref_p, smp.next_2d(), emitter_val);
    if (ds.pdf == 0.f) return mk3(0.f);
    Ray ray = spawn_ray_to(ref_p, ref_n, ds.p);
    float max_dist = ray.maxt;
    if (ref_si && mesh_is_transition(sc, ref_si->mesh)) medium = target_medium(sc, *ref_si, ray.d);
    float total_dist = 0.f; SurfaceInteraction si = invalid_si(); bool needs_intersection = true, active = true;
    while (active) {
        float remaining_dist = max_dist - total_dist;
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
                float t = fminf(remaining_dist, fminf(mei.t, si.t)) - mei.mint;
                V3 tr = exp3(mei.combined_extinction * (-t));
                V3 ffp = (si.t < mei.t || mei.t > remaining_dist) ? tr : tr * mei.combined_extinction;
                float tr_pdf = idx3(ffp, channel);
                transmittance *= tr_pdf > 0.f ? tr / tr_pdf : mk3(0.f);
            }
            if (mei.t > remaining_dist && mei_valid(mei)) total_dist = ds.dist;
            if (mei.t > remaining_dist) mei.t = LR_INF;
            escaped_medium = !mei_valid(mei); active_medium = mei_valid(mei);
            if (active_medium) {
                total_dist += mei.t; ray.o = mei.p; si.t = si.t - mei.t;
                if (is_spectral) transmittance *= mei.sigma_n;
                if (not_spectral) transmittance *= mei.sigma_n / mei.combined_extinction;
            }
        }
        if (active_surface && needs_intersection) { si = intersect(sc, tris_gid, ray, cnt); needs_intersection = false; }
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
    return transmittance * emitter_val;
}

// ===================== volpath (volpath.cpp:93-396) =====================
LR_DEV LaneResult sample_volpath(const DScene &sc, const DTri *tris_gid, Sampler &smp, Ray ray, int initial_medium, Counters &cnt) {
    bool valid_ray = !sc.hide_emitters && sc.env >= 0;
    float eta = 1.f; V3 throughput = mk3(1.f), result = mk3(0.f);
    int medium = initial_medium; MediumInteraction mei; mei.t = LR_INF; mei.medium = -1;
    bool specular_chain = !sc.hide_emitters; uint32_t depth = 0;
    uint32_t channel = (uint32_t) fminf(smp.next_1d() * 3.f, 2.f);
    SurfaceInteraction si = invalid_si(); bool needs_intersection = true;
    V3 last_scatter_p = mk3(0.f); float last_scatter_direction_pdf = 1.f;
    bool active = true;
    while (active) {
        active &= any_nonzero(throughput);
        float q = fminf(hmax(throughput) * sqr(eta), .95f);
        bool perform_rr = depth > (uint32_t) sc.rr_depth;
        if (active) { float u = smp.next_1d(); active &= (u < q) || !perform_rr; }
        if (perform_rr) throughput *= 1.f / q;
        active &= depth < sc.max_depth;
        if (!active) break;

        bool active_medium = medium >= 0, active_surface = !active_medium;
        bool act_null_scatter = false, act_medium_scatter = false, escaped_medium = false, is_spectral = false, not_spectral = false;
        if (active_medium) {
            const DMedium &m = sc.media[medium];
            is_spectral = m.m.has_spectral_extinction != 0; not_spectral = !is_spectral;
            mei = sample_interaction<false>(sc, medium, ray, smp.next_1d(), channel, nullptr, cnt.collisions);
            if (m.m.kind == LR_MEDIUM_HOMOGENEOUS && mei_valid(mei)) ray.maxt = mei.t;
            if (needs_intersection) si = intersect(sc, tris_gid, ray, cnt);
            needs_intersection = false;
            if (si.t < mei.t) mei.t = LR_INF;
            if (is_spectral) {
                float t = fminf(mei.t, si.t) - mei.mint;
                V3 tr = exp3(mei.combined_extinction * (-t));
                V3 pdf = si.t < mei.t ? tr : tr * mei.combined_extinction;
                float tr_pdf = idx3(pdf, channel);
                throughput *= tr_pdf > 0.f ? tr / tr_pdf : mk3(0.f);
            }
            escaped_medium = !mei_valid(mei); active_medium = mei_valid(mei);
            if (active_medium) {
                float null_scatter_prob = hmean(mei.sigma_n / mei.combined_extinction);
                bool null_scatter = smp.next_1d() < null_scatter_prob;
                act_null_scatter = null_scatter; act_medium_scatter = !null_scatter;
                if (is_spectral && act_null_scatter) throughput *= mei.sigma_n / null_scatter_prob;
                if (act_medium_scatter) { depth += 1; last_scatter_p = mei.p; cnt.real_scatters++; }
            }
        }
        active &= depth < sc.max_depth;
        act_medium_scatter &= active;
        if (act_null_scatter) { ray.o = mei.p; si.t = si.t - mei.t; }

        if (act_medium_scatter) {
            if (is_spectral) throughput *= mei.sigma_s / hmean(mei.sigma_t / mei.combined_extinction);
            if (not_spectral) throughput *= mei.sigma_s / mei.sigma_t;
            const DMedium &m = sc.media[mei.medium]; Phase phase; phase.kind = m.m.phase; phase.g = m.m.phase_g;
            bool sample_emitters = m.m.sample_emitters != 0;
            valid_ray = true; specular_chain = !sample_emitters;
            if (sample_emitters) {
                DirectionSample ds;
                V3 emitted = sample_emitter_vol(sc, tris_gid, mei.p, mk3(0.f), nullptr, smp, medium, channel, ds, cnt);
                float phase_val = phase.eval_pdf(mei.wi, ds.d);
                result += throughput * phase_val * emitted * mis_weight(ds.pdf, ds.delta ? 0.f : phase_val);
            }
            (void) smp.next_1d(); V2 s2 = smp.next_2d();
            float phase_pdf; V3 wo = phase.sample(mei.wi, s2, phase_pdf);
            if (phase_pdf > 0.f) {
                ray.o = mei.p; ray.d = wo; ray.maxt = LR_LARGEST;
                needs_intersection = true; last_scatter_direction_pdf = phase_pdf;
            }
        }

        active_surface |= escaped_medium;
        bool intersect_now = active_surface && needs_intersection;
        if (intersect_now) si = intersect(sc, tris_gid, ray, cnt);
        if (sc.hide_emitters && intersect_now && depth == 0u && si_valid(si) && sc.meshes[si.mesh].emitter >= 0) {   // volpath.cpp:307-323
            Ray r2 = spawn_ray(si.p, si.n, ray.d); Hit pi = skip_area_emitters(sc, tris_gid, r2, cnt);
            si = compute_si(sc, r2, pi, tris_gid);
        }
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
                V3 emitted = sample_emitter_vol(sc, tris_gid, si.p, si.n, &si, smp, medium, channel, ds, cnt);
                V3 wo = si.sh.to_local(ds.d);
                float bsdf_pdf; V3 bsdf_val = bsdf.eval_pdf(si.wi, wo, bsdf_pdf);
                result += throughput * bsdf_val * mis_weight(ds.pdf, ds.delta ? 0.f : bsdf_pdf) * emitted;
            }
            float s1 = smp.next_1d(); V2 s2 = smp.next_2d();
            V3 bsdf_val; BSDFSample bs = bsdf.sample(si.wi, s1, s2, bsdf_val);
            throughput *= bsdf_val; eta *= bs.eta;
            ray = spawn_ray(si.p, si.n, si.sh.to_world(bs.wo));
            needs_intersection = true;
            bool non_null = !(bs.sampled_type & F_NULL);
            if (non_null) { depth += 1; last_scatter_p = si.p; last_scatter_direction_pdf = bs.pdf; }
            valid_ray |= non_null;
            specular_chain |= non_null && (bs.sampled_type & F_DELTA);
            specular_chain &= !(bs.sampled_type & F_SMOOTH);
            if (mesh_is_transition(sc, si.mesh)) medium = target_medium(sc, si, ray.d);
        }
        active &= (active_surface || active_medium);
    }
    LaneResult out; out.L = result; out.valid = valid_ray; return out;
}

} // namespace lr
