/*
 * TEST INFRASTRUCTURE ONLY (see orc_math.h).  Scalar, one-lane-at-a-time
 * restatement of the reference integrators with the masking semantics of the
 * JIT variants (llvm_ad_rgb is the stated parity target): a lane's sampler only
 * advances where the reference draws with that lane's mask set.
 */
#pragma once
#include "orc_scene.h"
#include "orc_bsdf.h"
#include "orc_emitter.h"
#include "orc_film.h"

namespace orc {

struct Counters { uint64_t rays = 0, collisions = 0, real_scatters = 0; };

inline float mis_weight(float a, float b) { a *= a; b *= b; float w = a / (a + b); return std::isfinite(w) ? w : 0.f; }
inline float index_spectrum(V3 s, uint32_t ch) { return ch == 1 ? s.y : (ch == 2 ? s.z : s.x); }

/* PerspectiveCamera::sample_ray_differential (src/sensors/perspective.cpp:239-279) */
inline Ray camera_ray(const Scene &sc, V2 pos_sample) {
    V3 near_p = xform_point_proj(sc.sample_to_camera, V3(pos_sample.x, pos_sample.y, 0.f));
    V3 d = normalize(near_p);
    Ray r;
    r.o = V3(sc.cam_to_world.m[3], sc.cam_to_world.m[7], sc.cam_to_world.m[11]);
    r.d = xform_vector(sc.cam_to_world, d);
    float inv_z = 1.f / d.z, near_t = sc.sensor.near_clip * inv_z, far_t = sc.sensor.far_clip * inv_z;
    r.o = r.o + r.d * near_t;
    r.maxt = far_t - near_t;
    return r;
}

struct LaneResult { V3 L; bool valid; };

struct Integrator {
    const Scene &sc; EmitterOps em; Counters cnt;
    uint32_t max_depth; int rr_depth; bool hide_emitters;
    explicit Integrator(const Scene &s) : sc(s), em(s) {
        max_depth = s.integ.max_depth < 0 ? 0xffffffffu : (uint32_t) s.integ.max_depth;
        rr_depth = s.integ.rr_depth; hide_emitters = s.integ.hide_emitters != 0;
    }
    SurfaceInteraction intersect(const Ray &r) { cnt.rays++; return sc.ray_intersect(r); }
    /* Integrator::skip_area_emitters (src/render/integrator.cpp:97-123): first hit along `ray` that is not an
       area emitter; `ray` ends up as the segment that hit was found with. */
    Hit skip_area_emitters(Ray &ray) {
        for (;;) {
            cnt.rays++; Hit pi = sc.ray_intersect_prelim(ray);
            if (!(pi.valid() && sc.meshes[sc.tris[pi.prim].mesh].emitter >= 0)) return pi;
            SurfaceInteraction si = sc.compute_si(ray, pi);
            ray = si.spawn_ray(ray.d);
        }
    }

    /* ===================== path (src/integrators/path.cpp:95-349) ===================== */
    LaneResult sample_path(Sampler &smp, Ray ray) {
        if (max_depth == 0) return {V3(0.f), false};
        V3 throughput(1.f), result(0.f); float eta = 1.f; uint32_t depth = 0;
        bool valid_ray = !hide_emitters && sc.env >= 0;
        Interaction prev_si; float prev_bsdf_pdf = 1.f; bool prev_bsdf_delta = true;
        cnt.rays++; Hit pi = sc.ray_intersect_prelim(ray);
        if (hide_emitters && pi.valid() && sc.meshes[sc.tris[pi.prim].mesh].emitter >= 0) {   /* :178-191; ls.ray itself is kept */
            Ray r2 = sc.compute_si(ray, pi).spawn_ray(ray.d);
            pi = skip_area_emitters(r2);
        }
        bool active = true;
        while (active) {
            SurfaceInteraction si = sc.compute_si(ray, pi);
            int emitter = sc.emitter_of(si);
            if (emitter >= 0) {                                             /* :204-224 */
                DirectionSample ds = em.ds_from_si(si, prev_si);
                float em_pdf = 0.f;
                if (!prev_bsdf_delta) em_pdf = em.pdf_emitter_direction(prev_si, ds);
                float mis_bsdf = mis_weight(prev_bsdf_pdf, em_pdf);
                V3 ev = prev_bsdf_pdf > 0.f ? em.eval(emitter, si) : V3(0.f);
                result = fma3(throughput, ev * mis_bsdf, result);
            }
            bool active_next = (depth + 1 < max_depth) && si.is_valid();       /* :227 */
            if (!si.is_valid()) break;      /* JIT: masked lane, no further side effects */
            BSDF bsdf(sc, si);
            bool active_em = active_next && (bsdf.flags() & F_SMOOTH);
            DirectionSample ds; V3 em_weight(0.f), wo(0.f);
            V2 u_em = smp.next_2d();                                         /* :247-248, drawn by every loop-active lane */
            if (active_em) {
                ds = em.sample_emitter_direction(si, u_em, true, em_weight);
                if (ds.pdf != 0.f) cnt.rays++;
                active_em &= ds.pdf != 0.f;
                wo = si.to_local(ds.d);
            }
            float s1 = smp.next_1d(); V2 s2 = smp.next_2d();                 /* :268-269 */
            V3 bsdf_val = bsdf.eval(si.wi, wo); float bsdf_pdf = bsdf.pdf(si.wi, wo);   /* eval_pdf_sample = eval_pdf + sample (src/render/bsdf.cpp:21-31) */
            V3 bsdf_weight; BSDFSample bs = bsdf.sample(si.wi, s1, s2, bsdf_weight);
            if (active_em) {
                float mis_em = ds.delta ? 1.f : mis_weight(ds.pdf, bsdf_pdf);
                result = fma3(throughput, bsdf_val * em_weight * mis_em, result);
            }
            ray = si.spawn_ray(si.to_world(bs.wo));
            throughput *= bsdf_weight; eta *= bs.eta;
            valid_ray |= !(bs.sampled_type & F_NULL);
            prev_si = si; prev_bsdf_pdf = bs.pdf; prev_bsdf_delta = (bs.sampled_type & F_DELTA) != 0;
            depth += 1;
            float tmax = hmax(throughput), rr_prob = std::min(tmax * sqr(eta), .95f);
            bool rr_active = (int) depth >= rr_depth, rr_continue = smp.next_1d() < rr_prob;   /* :326-330 */
            if (rr_active) throughput *= 1.f / rr_prob;
            active = active_next && (!rr_active || rr_continue) && (tmax != 0.f);
            if (active) { cnt.rays++; pi = sc.ray_intersect_prelim(ray); }
        }
        return {valid_ray ? result : V3(0.f), valid_ray};
    }

    /* ===================== medium helpers ===================== */
    struct MediumInteraction : Interaction {
        V3 wi, sigma_s, sigma_n, sigma_t, combined_extinction, transmittance; float mint = 0.f; int medium = -1;
        uint32_t cidx[8]; float cw[8]; float grid_val = 0.f;   /* trilinear footprint (adjoint) */
        Ray spawn_ray_m(V3 d) const { return {p, d, Largest}; }  /* n = 0 => no offset (interaction.h:165-169) */
    };
    /* BoundingBox3f::ray_intersect (include/mitsuba/core/bbox.h:303-340) */
    static bool aabb_intersect(V3 bmin, V3 bmax, const Ray &ray, float &mint, float &maxt) {
        bool active = ray.d.x != 0.f || ray.d.y != 0.f || ray.d.z != 0.f;
        V3 d_rcp(1.f / ray.d.x, 1.f / ray.d.y, 1.f / ray.d.z), t_min, t_max;
        for (int k = 0; k < 3; ++k) {
            t_min[k] = ((d_rcp[k] >= 0 ? bmin[k] : bmax[k]) - ray.o[k]) * d_rcp[k];
            t_max[k] = ((d_rcp[k] >= 0 ? bmax[k] : bmin[k]) - ray.o[k]) * d_rcp[k];
        }
        auto maximum_safe = [](float a, float b) { return (a > b || !std::isfinite(b)) ? a : b; };
        auto minimum_safe = [](float a, float b) { return (a < b || !std::isfinite(b)) ? a : b; };
        active = active && !((t_min.x > t_max.y) || (t_min.y > t_max.x));
        t_min.x = maximum_safe(t_min.x, t_min.y); t_max.x = minimum_safe(t_max.x, t_max.y);
        active = active && !((t_min.x > t_max.z) || (t_min.z > t_max.x));
        t_min.x = maximum_safe(t_min.x, t_min.z); t_max.x = minimum_safe(t_max.x, t_max.z);
        mint = t_min.x; maxt = t_max.x; return active;
    }
    V3 majorant(const Medium &m) const {
        if (m.m.kind == LR_MEDIUM_HETEROGENEOUS) return V3(m.max_density);
        return V3(m.m.sigma_t_const[0], m.m.sigma_t_const[1], m.m.sigma_t_const[2]) * m.m.scale;
    }
    /* Medium::sample_interaction (src/render/medium.cpp:42-82) with
       heterogeneous.cpp:181-201 / homogeneous.cpp:160-181 coefficients */
    MediumInteraction sample_interaction(int mid, const Ray &ray, float sample, uint32_t channel) {
        const Medium &m = sc.media[mid]; MediumInteraction mei; mei.wi = -ray.d; mei.medium = mid; mei.n = V3(0.f);
        bool active = true; float mint = 0.f, maxt = Inf;
        if (m.m.kind == LR_MEDIUM_HETEROGENEOUS) {
            const Grid &g = sc.grids[m.m.sigma_t_grid];
            bool its = aabb_intersect(g.bmin, g.bmax, ray, mint, maxt);
            its &= (std::isfinite(mint) || std::isfinite(maxt));
            active = its;
            if (!active) { mint = 0.f; maxt = Inf; }
        }
        mint = std::max(0.f, mint); maxt = std::min(ray.maxt, maxt);
        V3 ce = majorant(m); float mj = index_spectrum(ce, channel);
        float sampled_t = mint + (-logf(1.f - sample) / mj);
        bool valid_mi = active && (sampled_t <= maxt);
        mei.t = valid_mi ? sampled_t : Inf;
        mei.p = ray_at(ray, sampled_t); mei.mint = mint;
        mei.sigma_s = mei.sigma_n = mei.sigma_t = V3(0.f);
        if (valid_mi) {
            V3 albedo = m.m.albedo_grid >= 0 ? sc.grids[m.m.albedo_grid].eval3(mei.p)
                                             : V3(m.m.albedo_const[0], m.m.albedo_const[1], m.m.albedo_const[2]);
            if (m.m.kind == LR_MEDIUM_HETEROGENEOUS) {
                cnt.collisions++;
                const Grid &g = sc.grids[m.m.sigma_t_grid]; float o[3];
                g.eval(mei.p, o, mei.cidx, mei.cw); mei.grid_val = o[0];
                V3 st = (g.ch == 1 ? V3(o[0]) : V3(o[0], o[1], o[2])) * m.m.scale;
                mei.sigma_t = st; mei.sigma_s = st * albedo; mei.sigma_n = V3(m.max_density) - st;
            } else {
                cnt.collisions++;
                V3 st = V3(m.m.sigma_t_const[0], m.m.sigma_t_const[1], m.m.sigma_t_const[2]) * m.m.scale;
                mei.sigma_t = st; mei.sigma_s = st * albedo; mei.sigma_n = V3(0.f);
            }
        }
        mei.combined_extinction = (m.m.kind == LR_MEDIUM_HOMOGENEOUS && !active) ? V3(0.f) : ce;
        return mei;
    }
    static V3 exp3(V3 a) { return {expf(a.x), expf(a.y), expf(a.z)}; }
    Phase phase_of(int mid) const { return Phase{sc.media[mid].m.phase, sc.media[mid].m.phase_g}; }

#include "orc_volpath.inl"
#include "orc_prbvolpath.inl"
#include "orc_biovolpath.inl"
#include "orc_volpathmis.inl"
};

} // namespace orc
