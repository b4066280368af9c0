/*
 * TEST INFRASTRUCTURE ONLY (see orc_math.h).  BSDFs, microfacet distribution,
 * Fresnel and phase functions restated from the reference.
 */
#pragma once
#include "orc_scene.h"

namespace orc {

enum : uint32_t {
    F_NULL = 1u, F_DIFFUSE_REFL = 2u, F_GLOSSY_REFL = 4u, F_GLOSSY_TRANS = 8u,
    F_DELTA_REFL = 16u, F_DELTA_TRANS = 32u,
    F_DELTA = F_NULL | F_DELTA_REFL | F_DELTA_TRANS,
    F_SMOOTH = F_DIFFUSE_REFL | F_GLOSSY_REFL | F_GLOSSY_TRANS
};

struct BSDFSample { V3 wo; float pdf = 0.f, eta = 0.f; uint32_t sampled_type = 0; };

/* fresnel (include/mitsuba/render/fresnel.h:35-75) */
inline void fresnel(float cos_theta_i, float eta, float &r, float &cos_theta_t, float &eta_it, float &eta_ti) {
    bool outside = cos_theta_i >= 0.f;
    float rcp_eta = 1.f / eta;
    eta_it = outside ? eta : rcp_eta; eta_ti = outside ? rcp_eta : eta;
    float cos_theta_t_sqr = fmaf(-fmaf(-cos_theta_i, cos_theta_i, 1.f), eta_ti * eta_ti, 1.f);
    float ci = fabsf(cos_theta_i), ct = safe_sqrt(cos_theta_t_sqr);
    bool index_matched = eta == 1.f, special = index_matched || ci == 0.f;
    float r_sc = index_matched ? 0.f : 1.f;
    float a_s = fmaf(-eta_it, ct, ci) / fmaf(eta_it, ct, ci);
    float a_p = fmaf(-eta_it, ci, ct) / fmaf(eta_it, ci, ct);
    r = 0.5f * (sqr(a_s) + sqr(a_p));
    if (special) r = r_sc;
    cos_theta_t = mulsign_neg(ct, cos_theta_i);
}
/* fresnel.h:276-320 */
inline V3 reflect(V3 wi) { return {-wi.x, -wi.y, wi.z}; }
inline V3 reflect(V3 wi, V3 m) { float s = 2.f * dot(wi, m); return {fmaf(m.x, s, -wi.x), fmaf(m.y, s, -wi.y), fmaf(m.z, s, -wi.z)}; }
inline V3 refract(V3 wi, float cos_theta_t, float eta_ti) { return {-eta_ti * wi.x, -eta_ti * wi.y, cos_theta_t}; }
inline V3 refract(V3 wi, V3 m, float cos_theta_t, float eta_ti) {
    float s = fmaf(dot(wi, m), eta_ti, cos_theta_t);
    return {fmaf(m.x, s, -(wi.x * eta_ti)), fmaf(m.y, s, -(wi.y * eta_ti)), fmaf(m.z, s, -(wi.z * eta_ti))};
}

/* MicrofacetDistribution (include/mitsuba/render/microfacet.h:185-423) */
struct Microfacet {
    uint32_t type; float au, av; bool sample_visible;
    Microfacet(uint32_t t, float a_u, float a_v, bool sv = true) : type(t), au(std::max(a_u, 1e-4f)), av(std::max(a_v, 1e-4f)), sample_visible(sv) {}
    void scale_alpha(float v) { au *= v; av *= v; }
    float eval(V3 m) const {
        float alpha_uv = au * av, ct = m.z, ct2 = sqr(ct), result;
        if (type == LR_MF_BECKMANN)
            result = expf(-(sqr(m.x / au) + sqr(m.y / av)) / ct2) / (Pi * alpha_uv * sqr(ct2));
        else
            result = 1.f / (Pi * alpha_uv * sqr(sqr(m.x / au) + sqr(m.y / av) + sqr(m.z)));
        return result * ct > 1e-20f ? result : 0.f;
    }
    float smith_g1(V3 v, V3 m) const {
        float xy_alpha_2 = sqr(au * v.x) + sqr(av * v.y), tan_theta_alpha_2 = xy_alpha_2 / sqr(v.z), result;
        if (type == LR_MF_BECKMANN) {
            float a = 1.f / sqrtf(tan_theta_alpha_2), a_sqr = sqr(a);
            result = a >= 1.6f ? 1.f : (3.535f * a + 2.181f * a_sqr) / (1.f + 2.276f * a + 2.577f * a_sqr);
        } else result = 2.f / (1.f + sqrtf(1.f + tan_theta_alpha_2));
        if (xy_alpha_2 == 0.f) result = 1.f;
        if (dot(v, m) * v.z <= 0.f) result = 0.f;
        return result;
    }
    float G(V3 wi, V3 wo, V3 m) const { return smith_g1(wi, m) * smith_g1(wo, m); }
    float pdf(V3 wi, V3 m) const {
        float result = eval(m);
        if (sample_visible) result *= smith_g1(wi, m) * abs_dot(wi, m) / wi.z; else result *= m.z;
        return result;
    }
    V2 sample_visible_11(float cos_theta_i, V2 sample) const {
        if (type == LR_MF_BECKMANN) {
            float tan_theta_i = safe_sqrt(fmaf(-cos_theta_i, cos_theta_i, 1.f)) / cos_theta_i;
            float cot_theta_i = 1.f / tan_theta_i;
            float maxval = erff(cot_theta_i);
            sample.x = std::max(std::min(sample.x, 1.f - 1e-6f), 1e-6f);
            sample.y = std::max(std::min(sample.y, 1.f - 1e-6f), 1e-6f);
            float x = maxval - (maxval + 1.f) * erff(sqrtf(-logf(sample.x)));
            sample.x *= 1.f + maxval + InvSqrtPi * tan_theta_i * expf(-sqr(cot_theta_i));
            for (int i = 0; i < 3; ++i) {
                float slope = erfinv_(x),
                      value = 1.f + x + InvSqrtPi * tan_theta_i * expf(-sqr(slope)) - sample.x,
                      derivative = 1.f - slope * tan_theta_i;
                x -= value / derivative;
            }
            return {erfinv_(x), erfinv_(fmaf(2.f, sample.y, -1.f))};
        } else {
            V2 p = square_to_uniform_disk_concentric(sample);
            float s = 0.5f * (1.f + cos_theta_i);
            p.y = lerp(safe_sqrt(1.f - sqr(p.x)), p.y, s);
            float x = p.x, y = p.y, z = safe_sqrt(1.f - fmaf(p.y, p.y, p.x * p.x));
            float sin_theta_i = safe_sqrt(1.f - sqr(cos_theta_i));
            float nrm = 1.f / fmaf(sin_theta_i, y, cos_theta_i * z);
            return {fmaf(cos_theta_i, y, -(sin_theta_i * z)) * nrm, x * nrm};
        }
    }
    void sample(V3 wi, V2 u, V3 &m, float &pdf_) const {
        if (!sample_visible) {
            float sin_phi, cos_phi, cos_theta, cos_theta_2, alpha_2;
            if (au == av) { sincos_(2.f * Pi * u.y, sin_phi, cos_phi); alpha_2 = au * au; }
            else {
                float ratio = av / au, tmp = ratio * tanf(2.f * Pi * u.y);
                cos_phi = 1.f / sqrtf(fmaf(tmp, tmp, 1.f));
                cos_phi = mulsign(cos_phi, fabsf(u.y - .5f) - .25f);
                sin_phi = cos_phi * tmp;
                alpha_2 = 1.f / (sqr(cos_phi / au) + sqr(sin_phi / av));
            }
            if (type == LR_MF_BECKMANN) {
                cos_theta = 1.f / sqrtf(fmaf(-alpha_2, logf(1.f - u.x), 1.f)); cos_theta_2 = sqr(cos_theta);
                float c3 = std::max(cos_theta_2 * cos_theta, 1e-20f);
                pdf_ = (1.f - u.x) / (Pi * au * av * c3);
            } else {
                float tan_theta_m_2 = alpha_2 * u.x / (1.f - u.x);
                cos_theta = 1.f / sqrtf(1.f + tan_theta_m_2); cos_theta_2 = sqr(cos_theta);
                float temp = 1.f + tan_theta_m_2 / alpha_2, c3 = std::max(cos_theta_2 * cos_theta, 1e-20f);
                pdf_ = 1.f / (Pi * au * av * c3 * sqr(temp));
            }
            float sin_theta = sqrtf(1.f - cos_theta_2);
            m = V3(cos_phi * sin_theta, sin_phi * sin_theta, cos_theta);
        } else {
            V3 wi_p = normalize(V3(au * wi.x, av * wi.y, wi.z));
            /* Frame3f::sincos_phi (include/mitsuba/core/frame.h) */
            float sin_theta_2 = fmaf(wi_p.x, wi_p.x, wi_p.y * wi_p.y), inv_sin_theta = 1.f / sqrtf(sin_theta_2);
            float sin_phi = wi_p.y * inv_sin_theta, cos_phi = wi_p.x * inv_sin_theta;
            if (fabsf(sin_theta_2) <= 4.f * Epsilon) { sin_phi = 0.f; cos_phi = 1.f; }
            sin_phi = std::min(std::max(sin_phi, -1.f), 1.f); cos_phi = std::min(std::max(cos_phi, -1.f), 1.f);
            float cos_theta = wi_p.z;
            V2 slope = sample_visible_11(cos_theta, u);
            slope = {fmaf(cos_phi, slope.x, -(sin_phi * slope.y)) * au, fmaf(sin_phi, slope.x, cos_phi * slope.y) * av};
            m = normalize(V3(-slope.x, -slope.y, 1.f));
            pdf_ = eval(m) * smith_g1(wi, m) * abs_dot(wi, m) / wi.z;
        }
    }
};

/* ---- principled helpers (src/bsdfs/principledhelpers.h) ------------------- */
struct GTR1 {
    float alpha;
    float eval(V3 m) const {
        float ct = m.z, ct2 = sqr(ct), a2 = sqr(alpha);
        float result = (a2 - 1.f) / (Pi * logf(a2) * (1.f + (a2 - 1.f) * ct2));
        return result * ct > 1e-20f ? result : 0.f;
    }
    float pdf(V3 m) const { return m.z < 0.f ? 0.f : m.z * eval(m); }
    V3 sample(V2 u) const {
        float sp, cp; sincos_(2.f * Pi * u.x, sp, cp);
        float a2 = sqr(alpha), ct2 = (1.f - powf(a2, 1.f - u.y)) / (1.f - a2);
        float st = sqrtf(std::max(0.f, 1.f - ct2)), ct = sqrtf(std::max(0.f, ct2));
        return {cp * st, sp * st, ct};
    }
};
inline float smith_ggx1(V3 v, V3 wh, float alpha) {
    float a2 = sqr(alpha), ct = fabsf(v.z), ct2 = sqr(ct), tan2 = (1.f - ct2) / ct2;
    float result = 2.f * (1.f / (1.f + sqrtf(1.f + a2 * tan2)));
    if (v.z == 1.f) result = 1.f;
    if (dot(v, wh) * v.z <= 0.f) result = 0.f;
    return result;
}
inline float clearcoat_G(V3 wi, V3 wo, V3 wh, float alpha) { return smith_ggx1(wi, wh, alpha) * smith_ggx1(wo, wh, alpha); }
inline float schlick_weight(float cos_i) { float m = std::min(std::max(1.f - cos_i, 0.f), 1.f); return sqr(sqr(m)) * m; }
inline float calc_schlick1(float R0, float cos_theta_i, float eta) {
    bool outside = cos_theta_i >= 0.f; float rcp_eta = 1.f / eta, eta_it = outside ? eta : rcp_eta, eta_ti = outside ? rcp_eta : eta;
    float ct2 = fmaf(-fmaf(-cos_theta_i, cos_theta_i, 1.f), sqr(eta_ti), 1.f), ct = safe_sqrt(ct2);
    return eta_it > 1.f ? lerp(schlick_weight(fabsf(cos_theta_i)), 1.f, R0) : lerp(schlick_weight(ct), 1.f, R0);
}
inline V3 calc_schlick3(V3 R0, float c, float eta) { return {calc_schlick1(R0.x, c, eta), calc_schlick1(R0.y, c, eta), calc_schlick1(R0.z, c, eta)}; }
inline float schlick_R0_eta(float eta) { return sqr((eta - 1.f) / (eta + 1.f)); }
inline bool mac_mic_compat(V3 m, V3 wi, V3 wo, float cos_theta_i, bool reflection) {
    if (reflection) return dot(wi, mulsign(m, cos_theta_i)) > 0.f && dot(wo, mulsign(m, cos_theta_i)) > 0.f;
    V3 mn(mulsign_neg(m.x, cos_theta_i), mulsign_neg(m.y, cos_theta_i), mulsign_neg(m.z, cos_theta_i));
    return dot(wi, mulsign(m, cos_theta_i)) > 0.f && dot(wo, mn) > 0.f;
}
inline float luminance(V3 c) { return c.x * 0.212671f + c.y * 0.715160f + c.z * 0.072169f; }
inline V3 principled_fresnel(float F_diel, float metallic, float spec_tint, V3 base, float lum, float cos_theta_i,
                             bool front_side, float bsdf, float eta, bool has_metallic, bool has_spec_tint) {
    bool outside = cos_theta_i >= 0.f; float rcp_eta = 1.f / eta, eta_it = outside ? eta : rcp_eta;
    V3 F_schlick(0.f);
    if (has_metallic) F_schlick += metallic * calc_schlick3(base, cos_theta_i, eta);
    if (has_spec_tint) {
        V3 c_tint = lum > 0.f ? base / lum : V3(1.f);
        V3 F0 = c_tint * schlick_R0_eta(eta_it);
        F_schlick += (1.f - metallic) * spec_tint * calc_schlick3(F0, cos_theta_i, eta);
    }
    V3 F_front = V3((1.f - metallic) * (1.f - spec_tint) * F_diel) + F_schlick;
    return front_side ? F_front : V3(bsdf * F_diel);
}
inline void calc_dist_params(float anisotropic, float roughness, bool has_aniso, float &ax, float &ay) {
    float r2 = sqr(roughness);
    if (!has_aniso) { ax = ay = std::max(0.001f, r2); return; }
    float aspect = sqrtf(1.f - 0.9f * anisotropic);
    ax = std::max(0.001f, r2 / aspect); ay = std::max(0.001f, r2 * aspect);
}
inline V3 sqrt3(V3 a) { return {sqrtf(a.x), sqrtf(a.y), sqrtf(a.z)}; }
inline V3 lerp3(V3 a, V3 b, float t) { return {lerp(a.x, b.x, t), lerp(a.y, b.y, t), lerp(a.z, b.z, t)}; }

/* ---- BSDF dispatch --------------------------------------------------------- */
struct BSDF {
    const lr_bsdf &b; V2 uv;
    const Scene *scene = nullptr;
    bool bump = false; Frame pf;        /* bumpmap: perturbed frame, relative to the original shading frame */
    explicit BSDF(const lr_bsdf &b_, V2 uv_ = {0.f, 0.f}) : b(b_), uv(uv_) {}
    /* the constructor the integrators use: resolves texture slots and the bump-mapped frame */
    BSDF(const Scene &sc, const SurfaceInteraction &si) : b(sc.bsdfs[sc.meshes[si.mesh].bsdf]), uv(si.uv), scene(&sc) {
        if (b.bump_tex) { bump = true; pf = bump_frame(sc.textures[b.bump_tex - 1], b, si); }
    }
    /* BumpMap::frame (src/bsdfs/bumpmap.cpp:248-276) */
    static Frame bump_frame(const Texture2D &tex, const lr_bsdf &b, const SurfaceInteraction &si) {
        V2 g = tex.eval_1_grad(si.uv); g.x *= b.bump_scale; g.y *= b.bump_scale;
        V3 dp_du = fma3(si.sh.n, g.x - dot(si.sh.n, si.dp_du), si.dp_du);
        V3 dp_dv = fma3(si.sh.n, g.y - dot(si.sh.n, si.dp_dv), si.dp_dv);
        Frame r;
        r.n = normalize(cross(dp_du, dp_dv));
        if (dot(si.n, r.n) < 0.f) r.n = -r.n;
        r.n = si.to_local(r.n);
        if (b.bump_flags & LR_BUMP_FLIP_INVALID_NORMALS) {
            if (si.wi.z * dot(si.wi, r.n) <= 0.f) r.n = V3(-r.n.x, -r.n.y, r.n.z);
        }
        /* Gram-Schmidt against the (world-space) dp_du, exactly as the reference writes it */
        r.s = normalize(fma3(r.n, -dot(r.n, si.dp_du), si.dp_du));
        r.t = cross(r.n, r.s);
        return r;
    }
    /* eval_shadow_terminator (src/bsdfs/normalmap_helpers.h:20-25; Frame::tan_theta_2, frame.h:80-83) */
    static float tan_theta_2(V3 v) { float t = fmaf(-v.z, v.z, 1.f); return std::max(t, 0.f) / sqr(v.z); }
    float shadow_terminator(V3 wo) const {
        float alpha2 = std::min(0.125f * tan_theta_2(pf.n), 1.f);
        return 2.f / (1.f + sqrtf(1.f + alpha2 * tan_theta_2(wo)));
    }
    /* BumpMap::sample / eval / pdf (bumpmap.cpp:136-201): nested BSDF in the perturbed frame */
    BSDFSample sample(V3 wi, float s1, V2 s2, V3 &weight) const {
        if (!bump) return sample_n(wi, s1, s2, weight);
        BSDFSample bs = sample_n(pf.to_local(wi), s1, s2, weight);
        bool active = weight.x != 0.f || weight.y != 0.f || weight.z != 0.f;
        V3 pwo = pf.to_world(bs.wo);
        active &= bs.wo.z * pwo.z > 0.f;
        bs.wo = pwo;
        if (b.bump_flags & LR_BUMP_SHADOWING_FUNCTION) weight = weight * shadow_terminator(bs.wo);
        if (!active) weight = V3(0.f);
        return bs;
    }
    V3 eval(V3 wi, V3 wo) const {
        if (!bump) return eval_n(wi, wo);
        V3 pwo = pf.to_local(wo);
        if (!(wo.z * pwo.z > 0.f)) return V3(0.f);
        V3 v = eval_n(pf.to_local(wi), pwo);
        if (b.bump_flags & LR_BUMP_SHADOWING_FUNCTION) v = v * shadow_terminator(wo);
        return v;
    }
    float pdf(V3 wi, V3 wo) const {
        if (!bump) return pdf_n(wi, wo);
        V3 pwo = pf.to_local(wo);
        if (!(wo.z * pwo.z > 0.f)) return 0.f;
        return pdf_n(pf.to_local(wi), pwo);
    }
    V3 refl() const {
        if (b.tex_kind == 2 && scene) return scene->textures[b.refl_tex - 1].eval(uv);   /* bitmap reflectance */
        if (b.tex_kind == 1) {                      /* checkerboard (src/textures/checkerboard.cpp:70-89) */
            float u = fmaf(b.tex_to_uv[0], uv.x, fmaf(b.tex_to_uv[1], uv.y, b.tex_to_uv[2])),
                  v = fmaf(b.tex_to_uv[3], uv.x, fmaf(b.tex_to_uv[4], uv.y, b.tex_to_uv[5]));
            bool mx = u - floorf(u) > .5f, my = v - floorf(v) > .5f;
            if (mx != my) return {b.tex_color1[0], b.tex_color1[1], b.tex_color1[2]};
        }
        return {b.reflectance[0], b.reflectance[1], b.reflectance[2]};
    }
    V3 trans() const { return {b.transmittance[0], b.transmittance[1], b.transmittance[2]}; }
    bool has(uint32_t f) const { return (b.flags & f) != 0; }

    uint32_t flags() const {
        switch (b.kind) {
            case LR_BSDF_DIFFUSE: return F_DIFFUSE_REFL;
            case LR_BSDF_ROUGHDIELECTRIC: return F_GLOSSY_REFL | F_GLOSSY_TRANS;
            case LR_BSDF_PRINCIPLED: return F_DIFFUSE_REFL | F_GLOSSY_REFL | (has(LR_PR_HAS_SPEC_TRANS) ? F_GLOSSY_TRANS : 0u);
            case LR_BSDF_NULL: return F_NULL;
            case LR_BSDF_DIELECTRIC: return F_DELTA_REFL | F_DELTA_TRANS;
        }
        return 0;
    }
    /* BSDF::eval_null_transmission (src/render/bsdf.cpp:33-36; null.cpp:80-86) */
    float eval_null_transmission() const { return b.kind == LR_BSDF_NULL ? 1.f : 0.f; }

    /* ---------------- sample ---------------- */
    BSDFSample sample_n(V3 wi, float s1, V2 s2, V3 &weight) const {
        BSDFSample bs; weight = V3(0.f);
        switch (b.kind) {
        case LR_BSDF_DIFFUSE: {                     /* src/bsdfs/diffuse.cpp:100-123 */
            if (!(wi.z > 0.f)) return bs;
            bs.wo = square_to_cosine_hemisphere(s2); bs.pdf = square_to_cosine_hemisphere_pdf(bs.wo);
            bs.eta = 1.f; bs.sampled_type = F_DIFFUSE_REFL;
            if (bs.pdf > 0.f) weight = refl();
            return bs;
        }
        case LR_BSDF_NULL: {                        /* src/bsdfs/null.cpp:41-67 */
            bs.wo = -wi; bs.sampled_type = F_NULL; bs.eta = 1.f; bs.pdf = 1.f; weight = V3(1.f); return bs;
        }
        case LR_BSDF_DIELECTRIC: {                  /* src/bsdfs/dielectric.cpp sample(), both lobes enabled */
            float r_i, ct, eta_it, eta_ti; fresnel(wi.z, b.eta, r_i, ct, eta_it, eta_ti);
            float t_i = 1.f - r_i; bool sel_r = s1 <= r_i;
            bs.pdf = sel_r ? r_i : t_i; bs.sampled_type = sel_r ? F_DELTA_REFL : F_DELTA_TRANS;
            bs.wo = sel_r ? reflect(wi) : refract(wi, ct, eta_ti); bs.eta = sel_r ? 1.f : eta_it;
            weight = sel_r ? refl() : trans() * sqr(eta_ti);
            return bs;
        }
        case LR_BSDF_ROUGHDIELECTRIC: {             /* src/bsdfs/roughdielectric.cpp:242-358 */
            float cos_theta_i = wi.z; bool active = cos_theta_i != 0.f;
            Microfacet distr(b.mf_kind, b.alpha_u, b.alpha_v, b.sample_visible != 0), sd(distr);
            if (!b.sample_visible) sd.scale_alpha(1.2f - .2f * sqrtf(fabsf(cos_theta_i)));
            V3 m; sd.sample(mulsign(wi, cos_theta_i), s2, m, bs.pdf);
            active &= bs.pdf != 0.f;
            float F, cos_theta_t, eta_it, eta_ti; fresnel(dot(wi, m), b.eta, F, cos_theta_t, eta_it, eta_ti);
            bool sel_r = s1 <= F && active, sel_t = !sel_r && active;
            V3 w(1.f);
            bs.pdf *= sel_r ? F : 1.f - F;
            bs.eta = sel_r ? 1.f : eta_it; bs.sampled_type = sel_r ? F_GLOSSY_REFL : F_GLOSSY_TRANS;
            float dwh_dwo = 0.f;
            if (sel_r) { bs.wo = reflect(wi, m); w *= refl(); dwh_dwo = 1.f / (4.f * dot(bs.wo, m)); }
            if (sel_t) {
                bs.wo = refract(wi, m, cos_theta_t, eta_ti);
                w *= trans() * sqr(eta_ti);
                dwh_dwo = (sqr(bs.eta) * dot(bs.wo, m)) / sqr(dot(wi, m) + bs.eta * dot(bs.wo, m));
            }
            if (b.sample_visible) w *= distr.smith_g1(bs.wo, m);
            else w *= distr.G(wi, bs.wo, m) * dot(wi, m) / (cos_theta_i * m.z);
            bs.pdf *= fabsf(dwh_dwo);
            weight = active ? w : V3(0.f);
            return bs;
        }
        case LR_BSDF_PRINCIPLED: return sample_principled(wi, s1, s2, weight);
        }
        return bs;
    }

    /* ---------------- eval / pdf ---------------- */
    V3 eval_n(V3 wi, V3 wo) const {
        switch (b.kind) {
        case LR_BSDF_DIFFUSE: {                     /* diffuse.cpp:125-141 */
            if (!(wi.z > 0.f && wo.z > 0.f)) return V3(0.f);
            return refl() * InvPi * wo.z;
        }
        case LR_BSDF_ROUGHDIELECTRIC: {             /* roughdielectric.cpp:360-434 */
            float ci = wi.z, co = wo.z; if (ci == 0.f) return V3(0.f);
            bool refl_ = ci * co > 0.f;
            float eta = ci > 0.f ? b.eta : 1.f / b.eta, inv_eta = ci > 0.f ? 1.f / b.eta : b.eta;
            V3 m = normalize(wi + wo * (refl_ ? 1.f : eta)); m = mulsign(m, m.z);
            Microfacet distr(b.mf_kind, b.alpha_u, b.alpha_v, b.sample_visible != 0);
            float D = distr.eval(m), F, ct, eit, eti; fresnel(dot(wi, m), b.eta, F, ct, eit, eti);
            float G = distr.G(wi, wo, m);
            if (refl_) return refl() * (F * D * G / (4.f * fabsf(ci)));
            float scale = sqr(inv_eta);
            float v = fabsf((scale * (1.f - F) * D * G * eta * eta * dot(wi, m) * dot(wo, m)) / (ci * sqr(dot(wi, m) + eta * dot(wo, m))));
            return trans() * v;
        }
        case LR_BSDF_PRINCIPLED: return eval_principled(wi, wo);
        default: return V3(0.f);
        }
    }
    float pdf_n(V3 wi, V3 wo) const {
        switch (b.kind) {
        case LR_BSDF_DIFFUSE: return (wi.z > 0.f && wo.z > 0.f) ? square_to_cosine_hemisphere_pdf(wo) : 0.f;
        case LR_BSDF_ROUGHDIELECTRIC: {             /* roughdielectric.cpp:436-499 */
            float ci = wi.z, co = wo.z; bool active = ci != 0.f;
            bool refl_ = ci * co > 0.f;
            float eta = ci > 0.f ? b.eta : 1.f / b.eta;
            V3 m = normalize(wi + wo * (refl_ ? 1.f : eta)); m = mulsign(m, m.z);
            active &= dot(wi, m) * wi.z > 0.f && dot(wo, m) * wo.z > 0.f;
            float dwh_dwo = refl_ ? 1.f / (4.f * dot(wo, m)) : (eta * eta * dot(wo, m)) / sqr(dot(wi, m) + eta * dot(wo, m));
            Microfacet sd(b.mf_kind, b.alpha_u, b.alpha_v, b.sample_visible != 0);
            if (!b.sample_visible) sd.scale_alpha(1.2f - .2f * sqrtf(fabsf(wi.z)));
            float prob = sd.pdf(mulsign(wi, wi.z), m);
            float F, ct, eit, eti; fresnel(dot(wi, m), b.eta, F, ct, eit, eti);
            prob *= refl_ ? F : 1.f - F;
            return active ? prob * fabsf(dwh_dwo) : 0.f;
        }
        case LR_BSDF_PRINCIPLED: return pdf_principled(wi, wo);
        default: return 0.f;
        }
    }

    /* ---------------- principled (src/bsdfs/principled.cpp) ---------------- */
    struct PrParams { float anisotropic, roughness, spec_trans, metallic, clearcoat, flatness, sheen, brdf, bsdf; };
    PrParams pr() const {
        PrParams p;
        p.anisotropic = has(LR_PR_HAS_ANISOTROPIC) ? b.anisotropic : 0.f; p.roughness = b.roughness;
        p.spec_trans = has(LR_PR_HAS_SPEC_TRANS) ? b.spec_trans : 0.f; p.metallic = has(LR_PR_HAS_METALLIC) ? b.metallic : 0.f;
        p.clearcoat = has(LR_PR_HAS_CLEARCOAT) ? b.clearcoat : 0.f; p.flatness = has(LR_PR_HAS_FLATNESS) ? b.flatness : 0.f;
        p.sheen = has(LR_PR_HAS_SHEEN) ? b.sheen : 0.f;
        p.brdf = (1.f - p.metallic) * (1.f - p.spec_trans); p.bsdf = (1.f - p.metallic) * p.spec_trans;
        return p;
    }
    void pr_probs(const PrParams &p, bool front, float F, float &ps_r, float &ps_t, float &p_cc, float &p_d) const {
        ps_r = front ? b.main_spec_srate * (1.f - p.bsdf * (1.f - F)) : F;
        ps_t = has(LR_PR_HAS_SPEC_TRANS) ? (front ? b.main_spec_srate * p.bsdf * (1.f - F) : (1.f - F)) : 0.f;
        p_cc = has(LR_PR_HAS_CLEARCOAT) ? (front ? 0.25f * p.clearcoat * b.clearcoat_srate : 0.f) : 0.f;
        p_d = front ? p.brdf * b.diff_refl_srate : 0.f;
    }
    /* principled.cpp:333-491 */
    BSDFSample sample_principled(V3 wi, float s1, V2 s2, V3 &weight) const {
        BSDFSample bs; weight = V3(0.f);
        float cos_theta_i = wi.z; bool active = cos_theta_i != 0.f;
        if (!active) return bs;
        PrParams p = pr();
        float bsdf_s = has(LR_PR_HAS_SPEC_TRANS) ? p.bsdf : 0.f;   /* sample(): `bsdf` is 0 without spec_trans (:349-350) */
        bool front = cos_theta_i > 0.f;
        float ax, ay; calc_dist_params(p.anisotropic, p.roughness, has(LR_PR_HAS_ANISOTROPIC), ax, ay);
        Microfacet sd(LR_MF_GGX, ax, ay);
        V3 m_spec; float pdf_m; sd.sample(mulsign(wi, cos_theta_i), s2, m_spec, pdf_m);
        float F, cos_theta_t, eta_it, eta_ti; fresnel(dot(wi, m_spec), b.eta, F, cos_theta_t, eta_it, eta_ti);
        active &= (front || bsdf_s > 0.f);
        PrParams ps = p; ps.bsdf = bsdf_s;
        float ps_r, ps_t, p_cc, p_d; pr_probs(ps, front, F, ps_r, ps_t, p_cc, p_d);
        float rcp_tot = 1.f / (ps_r + ps_t + p_cc + p_d);
        ps_t *= rcp_tot; p_cc *= rcp_tot; p_d *= rcp_tot;
        float curr = 0.f;
        bool s_diff = active && s1 < p_d; curr += p_d;
        bool s_cc = has(LR_PR_HAS_CLEARCOAT) && active && s1 >= curr && s1 < curr + p_cc; curr += p_cc;
        bool s_tr = has(LR_PR_HAS_SPEC_TRANS) && active && s1 >= curr && s1 < curr + ps_t; curr += ps_t;
        bool s_rf = active && s1 >= curr;
        bs.eta = 1.f;
        if (s_rf) {
            V3 wo = reflect(wi, m_spec); bs.wo = wo; bs.sampled_type = F_GLOSSY_REFL;
            bool refl_ = cos_theta_i * wo.z > 0.f;
            active &= mac_mic_compat(m_spec, wi, wo, cos_theta_i, true) && refl_;
        }
        if (s_tr) {
            V3 wo = refract(wi, m_spec, cos_theta_t, eta_ti); bs.wo = wo; bs.sampled_type = F_GLOSSY_TRANS; bs.eta = eta_it;
            bool refr = cos_theta_i * wo.z < 0.f;
            active &= mac_mic_compat(m_spec, wi, wo, cos_theta_i, false) && refr;
        }
        if (s_cc) {
            GTR1 cc{lerp(0.1f, 0.001f, b.clearcoat_gloss)};
            V3 m_cc = cc.sample(s2); V3 wo = reflect(wi, m_cc); bs.wo = wo; bs.sampled_type = F_GLOSSY_REFL;
            bool refl_ = cos_theta_i * wo.z > 0.f;
            active &= mac_mic_compat(m_cc, wi, wo, cos_theta_i, true) && refl_;
        }
        if (s_diff) {
            V3 wo = square_to_cosine_hemisphere(s2); bs.wo = wo; bs.sampled_type = F_DIFFUSE_REFL;
            active &= cos_theta_i * wo.z > 0.f;
        }
        bs.pdf = active ? pdf_principled(wi, bs.wo) : 0.f;
        active &= bs.pdf > 0.f;
        if (active) weight = eval_principled(wi, bs.wo) / bs.pdf;
        return bs;
    }
    /* principled.cpp:493-709 */
    V3 eval_principled(V3 wi, V3 wo) const {
        float cos_theta_i = wi.z; if (cos_theta_i == 0.f) return V3(0.f);
        PrParams p = pr(); V3 base = refl();
        float cos_theta_o = wo.z;
        bool refl_ = cos_theta_i * cos_theta_o > 0.f, refr = cos_theta_i * cos_theta_o < 0.f, front = cos_theta_i > 0.f;
        float inv_eta = 1.f / b.eta, eta_path = front ? b.eta : inv_eta, inv_eta_path = front ? inv_eta : b.eta;
        float ax, ay; calc_dist_params(p.anisotropic, p.roughness, has(LR_PR_HAS_ANISOTROPIC), ax, ay);
        Microfacet sd(LR_MF_GGX, ax, ay);
        V3 wh = normalize(wi + wo * (refl_ ? 1.f : eta_path)); wh = mulsign(wh, wh.z);
        float F, ct, eit, eti; fresnel(dot(wi, wh), b.eta, F, ct, eit, eti);
        bool rc = mac_mic_compat(wh, wi, wo, cos_theta_i, true), tc = mac_mic_compat(wh, wi, wo, cos_theta_i, false);
        bool a_sr = refl_ && rc && F > 0.f;
        bool a_cc = has(LR_PR_HAS_CLEARCOAT) && p.clearcoat > 0.f && refl_ && rc && front;
        bool a_st = has(LR_PR_HAS_SPEC_TRANS) && p.bsdf > 0.f && refr && tc && F < 1.f;
        bool a_d = p.brdf > 0.f && refl_ && front;
        bool a_sh = has(LR_PR_HAS_SHEEN) && p.sheen > 0.f && refl_ && (1.f - p.metallic > 0.f) && front;
        float D = sd.eval(wh), G = sd.G(wi, wo, wh);
        V3 value(0.f);
        if (a_sr) {
            float lum = has(LR_PR_HAS_SPEC_TINT) ? luminance(base) : 1.f;
            float spec_tint = has(LR_PR_HAS_SPEC_TINT) ? b.spec_tint : 0.f;
            V3 Fp = principled_fresnel(F, p.metallic, spec_tint, base, lum, dot(wi, wh), front, p.bsdf, b.eta,
                                       has(LR_PR_HAS_METALLIC), has(LR_PR_HAS_SPEC_TINT));
            value += Fp * D * G / (4.f * fabsf(cos_theta_i));
        }
        if (a_st) {
            float scale = sqr(inv_eta_path);
            value += sqrt3(base) * p.bsdf *
                     fabsf((scale * (1.f - F) * D * G * eta_path * eta_path * dot(wi, wh) * dot(wo, wh)) /
                           (cos_theta_i * sqr(dot(wi, wh) + eta_path * dot(wo, wh))));
        }
        if (a_cc) {
            float Fcc = calc_schlick1(0.04f, dot(wi, wh), b.eta);
            GTR1 mf{lerp(0.1f, 0.001f, b.clearcoat_gloss)};
            float Dcc = mf.eval(wh), Gcc = clearcoat_G(wi, wo, wh, 0.25f);
            value += V3((p.clearcoat * 0.25f) * Fcc * Dcc * Gcc * fabsf(cos_theta_o));
        }
        if (a_d) {
            float Fo = schlick_weight(fabsf(cos_theta_o)), Fi = schlick_weight(fabsf(cos_theta_i));
            float f_diff = (1.f - 0.5f * Fi) * (1.f - 0.5f * Fo);
            float cos_theta_d = dot(wh, wo), Rr = 2.f * p.roughness * sqr(cos_theta_d);
            float f_retro = Rr * (Fo + Fi + Fo * Fi * (Rr - 1.f));
            if (has(LR_PR_HAS_FLATNESS)) {
                float Fss90 = Rr / 2.f, Fss = lerp(1.f, Fss90, Fo) * lerp(1.f, Fss90, Fi);
                float f_ss = 1.25f * (Fss * (1.f / (fabsf(cos_theta_o) + fabsf(cos_theta_i)) - 0.5f) + 0.5f);
                value += p.brdf * fabsf(cos_theta_o) * base * InvPi * lerp(f_diff + f_retro, f_ss, p.flatness);
            } else
                value += p.brdf * fabsf(cos_theta_o) * base * InvPi * (f_diff + f_retro);
            if (a_sh) {
                float Fd = schlick_weight(fabsf(cos_theta_d));
                if (has(LR_PR_HAS_SHEEN_TINT)) {
                    float lum = luminance(base); V3 c_tint = lum > 0.f ? base / lum : V3(1.f);
                    V3 c_sheen = lerp3(V3(1.f), c_tint, b.sheen_tint);
                    value += p.sheen * (1.f - p.metallic) * Fd * c_sheen * fabsf(cos_theta_o);
                } else
                    value += V3(p.sheen * (1.f - p.metallic) * Fd * fabsf(cos_theta_o));
            }
        }
        return value;
    }
    /* principled.cpp:711-822 */
    float pdf_principled(V3 wi, V3 wo) const {
        float cos_theta_i = wi.z; if (cos_theta_i == 0.f) return 0.f;
        PrParams p = pr(); bool front = cos_theta_i > 0.f;
        float eta_path = front ? b.eta : 1.f / b.eta, cos_theta_o = wo.z;
        bool refl_ = cos_theta_i * cos_theta_o > 0.f, refr = cos_theta_i * cos_theta_o < 0.f;
        V3 wh = normalize(wi + wo * (refl_ ? 1.f : eta_path)); wh = mulsign(wh, wh.z);
        float ax, ay; calc_dist_params(p.anisotropic, p.roughness, has(LR_PR_HAS_ANISOTROPIC), ax, ay);
        Microfacet sd(LR_MF_GGX, ax, ay);
        float F, ct, eit, eti; fresnel(dot(wi, wh), b.eta, F, ct, eit, eti);
        float ps_r, ps_t, p_cc, p_d; pr_probs(p, front, F, ps_r, ps_t, p_cc, p_d);
        float rcp_tot = 1.f / (ps_r + ps_t + p_cc + p_d);
        ps_r *= rcp_tot; ps_t *= rcp_tot; p_cc *= rcp_tot; p_d *= rcp_tot;
        float dwh_dwo_abs;
        if (has(LR_PR_HAS_SPEC_TRANS)) {
            float dih = dot(wi, wh), doh = dot(wo, wh);
            dwh_dwo_abs = fabsf(refl_ ? 1.f / (4.f * doh) : (sqr(eta_path) * doh) / sqr(dih + eta_path * doh));
        } else dwh_dwo_abs = fabsf(1.f / (4.f * dot(wo, wh)));
        float pdf = 0.f;
        bool mr = mac_mic_compat(wh, wi, wo, cos_theta_i, true) && refl_;
        if (mr) pdf += ps_r * sd.pdf(mulsign(wi, cos_theta_i), wh) * dwh_dwo_abs;
        if (refl_) pdf += p_d * square_to_cosine_hemisphere_pdf(wo);
        if (has(LR_PR_HAS_SPEC_TRANS)) {
            bool mt = mac_mic_compat(wh, wi, wo, cos_theta_i, false) && refr;
            if (mt) pdf += ps_t * sd.pdf(mulsign(wi, cos_theta_i), wh) * dwh_dwo_abs;
        }
        if (has(LR_PR_HAS_CLEARCOAT)) {
            GTR1 cc{lerp(0.1f, 0.001f, b.clearcoat_gloss)};
            if (mr) pdf += p_cc * cc.pdf(wh) * dwh_dwo_abs;
        }
        return pdf;
    }
};

/* ---- phase functions -------------------------------------------------------- */
/* HG (src/phase/hg.cpp:64-99), isotropic (src/phase/isotropic.cpp:39-58).
   `wi` is the medium interaction's incident direction (= -ray.d), frame = Frame(wi). */
struct Phase {
    uint32_t kind; float g;
    float eval_hg(float cos_theta) const {
        float temp = 1.f + sqr(g) + 2.f * g * cos_theta;
        return InvFourPi * (1.f - sqr(g)) / (temp * sqrtf(temp));
    }
    V3 sample(V3 wi, V2 u, float &pdf) const {
        if (kind == LR_PHASE_ISOTROPIC) { pdf = InvFourPi; return square_to_uniform_sphere(u); }
        float sqr_term = (1.f - sqr(g)) / (1.f - g + 2.f * g * u.x), cos_theta = (1.f + sqr(g) - sqr(sqr_term)) / (2.f * g);
        if (fabsf(g) < Epsilon) cos_theta = 1.f - 2.f * u.x;
        float sin_theta = safe_sqrt(1.f - sqr(cos_theta)), sp, cp; sincos_(2.f * Pi * u.y, sp, cp);
        Frame f(wi);
        V3 wo = f.to_world(V3(sin_theta * cp, sin_theta * sp, -cos_theta));
        pdf = eval_hg(-cos_theta);
        return wo;
    }
    float eval_pdf(V3 wi, V3 wo) const {
        if (kind == LR_PHASE_ISOTROPIC) return InvFourPi;
        return eval_hg(dot(wo, wi));
    }
};

} // namespace orc
