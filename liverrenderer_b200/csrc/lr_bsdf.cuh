// lr_bsdf.cuh -- surface scattering models and phase functions on the hot path (device code).
// diffuse (src/bsdfs/diffuse.cpp:100-178), roughdielectric (src/bsdfs/roughdielectric.cpp:242-499),
// principled (src/bsdfs/principled.cpp:333-822 + principledhelpers.h), null (src/bsdfs/null.cpp:41-86),
// dielectric (src/bsdfs/dielectric.cpp), MicrofacetDistribution (include/mitsuba/render/microfacet.h:185-423),
// fresnel/reflect/refract (include/mitsuba/render/fresnel.h:35-75,276-320),
// hg (src/phase/hg.cpp:64-99), isotropic (src/phase/isotropic.cpp:39-58).
#pragma once
#include "lr_scene.cuh"

namespace lr {

enum : uint32_t {
    F_NULL = 1u, F_DIFFUSE_REFL = 2u, F_GLOSSY_REFL = 4u, F_GLOSSY_TRANS = 8u, F_DELTA_REFL = 16u, F_DELTA_TRANS = 32u,
    F_DELTA = F_NULL | F_DELTA_REFL | F_DELTA_TRANS, F_SMOOTH = F_DIFFUSE_REFL | F_GLOSSY_REFL | F_GLOSSY_TRANS
};
struct BSDFSample { V3 wo; float pdf, eta; uint32_t sampled_type; };

LR_DEV void fresnel(float cos_theta_i, float eta, float &r, float &cos_theta_t, float &eta_it, float &eta_ti) {
    bool outside = cos_theta_i >= 0.f;
    float rcp_eta = 1.f / eta;
    eta_it = outside ? eta : rcp_eta; eta_ti = outside ? rcp_eta : eta;
    float cos_theta_t_sqr = fmaf(-fmaf(-cos_theta_i, cos_theta_i, 1.f), __fmul_rn(eta_ti, eta_ti), 1.f);
    float ci = fabsf(cos_theta_i), ct = safe_sqrt(cos_theta_t_sqr);
    bool index_matched = eta == 1.f, special = index_matched || ci == 0.f;
    float a_s = fmaf(-eta_it, ct, ci) / fmaf(eta_it, ct, ci);
    float a_p = fmaf(-eta_it, ci, ct) / fmaf(eta_it, ci, ct);
    r = 0.5f * (sqr(a_s) + sqr(a_p));
    if (special) r = index_matched ? 0.f : 1.f;
    cos_theta_t = mulsign_neg(ct, cos_theta_i);
}
LR_DEV V3 reflect_l(V3 wi) { return mk3(-wi.x, -wi.y, wi.z); }
LR_DEV V3 reflect_m(V3 wi, V3 m) { float s = 2.f * dot(wi, m); return mk3(fmaf(m.x, s, -wi.x), fmaf(m.y, s, -wi.y), fmaf(m.z, s, -wi.z)); }
LR_DEV V3 refract_l(V3 wi, float cos_theta_t, float eta_ti) { return mk3(-eta_ti * wi.x, -eta_ti * wi.y, cos_theta_t); }
LR_DEV V3 refract_m(V3 wi, V3 m, float cos_theta_t, float eta_ti) {
    float s = fmaf(dot(wi, m), eta_ti, cos_theta_t);
    return mk3(fmaf(m.x, s, -__fmul_rn(wi.x, eta_ti)), fmaf(m.y, s, -__fmul_rn(wi.y, eta_ti)), fmaf(m.z, s, -__fmul_rn(wi.z, eta_ti)));
}

struct Microfacet {
    uint32_t type; float au, av; bool sample_visible;
    LR_DEV void init(uint32_t t, float a_u, float a_v, bool sv) { type = t; au = fmaxf(a_u, 1e-4f); av = fmaxf(a_v, 1e-4f); sample_visible = sv; }
    LR_DEV void scale_alpha(float v) { au *= v; av *= v; }
    LR_DEV float eval(V3 m) const {
        float alpha_uv = au * av, ct = m.z, ct2 = sqr(ct), result;
        if (type == LR_MF_BECKMANN) result = expf(-(sqr(m.x / au) + sqr(m.y / av)) / ct2) / (kPi * alpha_uv * sqr(ct2));
        else result = 1.f / (kPi * alpha_uv * sqr(sqr(m.x / au) + sqr(m.y / av) + sqr(m.z)));
        return result * ct > 1e-20f ? result : 0.f;
    }
    LR_DEV float smith_g1(V3 v, V3 m) const {
        float xy_alpha_2 = sqr(au * v.x) + sqr(av * v.y), tan_theta_alpha_2 = xy_alpha_2 / sqr(v.z), result;
        if (type == LR_MF_BECKMANN) {
            float a = 1.f / sqrtf(tan_theta_alpha_2), a_sqr = sqr(a);
            result = a >= 1.6f ? 1.f : (3.535f * a + 2.181f * a_sqr) / (1.f + 2.276f * a + 2.577f * a_sqr);
        } else result = 2.f / (1.f + sqrtf(1.f + tan_theta_alpha_2));
        if (xy_alpha_2 == 0.f) result = 1.f;
        if (dot(v, m) * v.z <= 0.f) result = 0.f;
        return result;
    }
    LR_DEV float G(V3 wi, V3 wo, V3 m) const { return smith_g1(wi, m) * smith_g1(wo, m); }
    LR_DEV float pdf(V3 wi, V3 m) const {
        float result = eval(m);
        if (sample_visible) result *= smith_g1(wi, m) * abs_dot(wi, m) / wi.z; else result *= m.z;
        return result;
    }
    LR_DEV V2 sample_visible_11(float cos_theta_i, V2 sample) const {
        if (type == LR_MF_BECKMANN) {
            float tan_theta_i = safe_sqrt(fmaf(-cos_theta_i, cos_theta_i, 1.f)) / cos_theta_i;
            float cot_theta_i = 1.f / tan_theta_i;
            float maxval = erff(cot_theta_i);
            sample.x = fmaxf(fminf(sample.x, 1.f - 1e-6f), 1e-6f); sample.y = fmaxf(fminf(sample.y, 1.f - 1e-6f), 1e-6f);
            float x = maxval - (maxval + 1.f) * erff(sqrtf(-logf(sample.x)));
            sample.x *= 1.f + maxval + kInvSqrtPi * tan_theta_i * expf(-sqr(cot_theta_i));
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                float slope = erfinv_giles(x), value = 1.f + x + kInvSqrtPi * tan_theta_i * expf(-sqr(slope)) - sample.x,
                      derivative = 1.f - slope * tan_theta_i;
                x -= value / derivative;
            }
            return mk2(erfinv_giles(x), erfinv_giles(fmaf(2.f, sample.y, -1.f)));
        } else {
            V2 p = square_to_uniform_disk_concentric(sample);
            float s = 0.5f * (1.f + cos_theta_i);
            p.y = lerpf(safe_sqrt(1.f - sqr(p.x)), p.y, s);
            float x = p.x, y = p.y, z = safe_sqrt(1.f - fmaf(p.y, p.y, __fmul_rn(p.x, p.x)));
            float sin_theta_i = safe_sqrt(1.f - sqr(cos_theta_i));
            float nrm = 1.f / fmaf(sin_theta_i, y, __fmul_rn(cos_theta_i, z));
            return mk2(fmaf(cos_theta_i, y, -__fmul_rn(sin_theta_i, z)) * nrm, x * nrm);
        }
    }
    LR_DEV void sample(V3 wi, V2 u, V3 &m, float &pdf_) const {
        if (!sample_visible) {
            float sin_phi, cos_phi, cos_theta, cos_theta_2, alpha_2;
            if (au == av) { sincosf(2.f * kPi * u.y, &sin_phi, &cos_phi); alpha_2 = au * au; }
            else {
                float ratio = av / au, tmp = ratio * tanf(2.f * kPi * u.y);
                cos_phi = 1.f / sqrtf(fmaf(tmp, tmp, 1.f)); cos_phi = mulsign(cos_phi, fabsf(u.y - .5f) - .25f);
                sin_phi = cos_phi * tmp; alpha_2 = 1.f / (sqr(cos_phi / au) + sqr(sin_phi / av));
            }
            if (type == LR_MF_BECKMANN) {
                cos_theta = 1.f / sqrtf(fmaf(-alpha_2, logf(1.f - u.x), 1.f)); cos_theta_2 = sqr(cos_theta);
                float c3 = fmaxf(cos_theta_2 * cos_theta, 1e-20f); pdf_ = (1.f - u.x) / (kPi * au * av * c3);
            } else {
                float tan_theta_m_2 = alpha_2 * u.x / (1.f - u.x);
                cos_theta = 1.f / sqrtf(1.f + tan_theta_m_2); cos_theta_2 = sqr(cos_theta);
                float temp = 1.f + tan_theta_m_2 / alpha_2, c3 = fmaxf(cos_theta_2 * cos_theta, 1e-20f);
                pdf_ = 1.f / (kPi * au * av * c3 * sqr(temp));
            }
            float sin_theta = sqrtf(1.f - cos_theta_2);
            m = mk3(cos_phi * sin_theta, sin_phi * sin_theta, cos_theta);
        } else {
            V3 wi_p = normalize(mk3(au * wi.x, av * wi.y, wi.z));
            float sin_theta_2 = fmaf(wi_p.x, wi_p.x, __fmul_rn(wi_p.y, wi_p.y)), inv_sin_theta = 1.f / sqrtf(sin_theta_2);   // Frame3f::sincos_phi
            float sin_phi = wi_p.y * inv_sin_theta, cos_phi = wi_p.x * inv_sin_theta;
            sin_phi = fminf(fmaxf(sin_phi, -1.f), 1.f); cos_phi = fminf(fmaxf(cos_phi, -1.f), 1.f);
            if (fabsf(sin_theta_2) <= 4.f * kEps) { sin_phi = 0.f; cos_phi = 1.f; }
            V2 slope = sample_visible_11(wi_p.z, u);
            slope = mk2(fmaf(cos_phi, slope.x, -__fmul_rn(sin_phi, slope.y)) * au, fmaf(sin_phi, slope.x, __fmul_rn(cos_phi, slope.y)) * av);
            m = normalize(mk3(-slope.x, -slope.y, 1.f));
            pdf_ = eval(m) * smith_g1(wi, m) * abs_dot(wi, m) / wi.z;
        }
    }
};

// ---- principled helpers (src/bsdfs/principledhelpers.h) --------------------------------
LR_DEV float gtr1_eval(float alpha, V3 m) {
    float ct = m.z, ct2 = sqr(ct), a2 = sqr(alpha);
    float result = (a2 - 1.f) / (kPi * logf(a2) * (1.f + (a2 - 1.f) * ct2));
    return result * ct > 1e-20f ? result : 0.f;
}
LR_DEV float gtr1_pdf(float alpha, V3 m) { return m.z < 0.f ? 0.f : m.z * gtr1_eval(alpha, m); }
LR_DEV V3 gtr1_sample(float alpha, V2 u) {
    float sp, cp; sincosf(2.f * kPi * u.x, &sp, &cp);
    float a2 = sqr(alpha), ct2 = (1.f - powf(a2, 1.f - u.y)) / (1.f - a2);
    float st = sqrtf(fmaxf(0.f, 1.f - ct2)), ct = sqrtf(fmaxf(0.f, ct2));
    return mk3(cp * st, sp * st, ct);
}
LR_DEV float smith_ggx1(V3 v, V3 wh, float alpha) {
    float a2 = sqr(alpha), ct = fabsf(v.z), ct2 = sqr(ct), tan2 = (1.f - ct2) / ct2;
    float result = 2.f * (1.f / (1.f + sqrtf(1.f + a2 * tan2)));
    if (v.z == 1.f) result = 1.f;
    if (dot(v, wh) * v.z <= 0.f) result = 0.f;
    return result;
}
LR_DEV float schlick_weight(float cos_i) { float m = fminf(fmaxf(1.f - cos_i, 0.f), 1.f); return sqr(sqr(m)) * m; }
LR_DEV float calc_schlick1(float R0, float cos_theta_i, float eta) {
    bool outside = cos_theta_i >= 0.f; float rcp_eta = 1.f / eta, eta_it = outside ? eta : rcp_eta, eta_ti = outside ? rcp_eta : eta;
    float ct2 = fmaf(-fmaf(-cos_theta_i, cos_theta_i, 1.f), sqr(eta_ti), 1.f), ct = safe_sqrt(ct2);
    return eta_it > 1.f ? lerpf(schlick_weight(fabsf(cos_theta_i)), 1.f, R0) : lerpf(schlick_weight(ct), 1.f, R0);
}
LR_DEV V3 calc_schlick3(V3 R0, float c, float eta) { return mk3(calc_schlick1(R0.x, c, eta), calc_schlick1(R0.y, c, eta), calc_schlick1(R0.z, c, eta)); }
LR_DEV bool mac_mic_compat(V3 m, V3 wi, V3 wo, float cos_theta_i, bool reflection) {
    V3 ms = mulsign(m, cos_theta_i);
    if (reflection) return dot(wi, ms) > 0.f && dot(wo, ms) > 0.f;
    return dot(wi, ms) > 0.f && dot(wo, -ms) > 0.f;
}
LR_DEV void calc_dist_params(float anisotropic, float roughness, bool has_aniso, float &ax, float &ay) {
    float r2 = sqr(roughness);
    if (!has_aniso) { ax = ay = fmaxf(0.001f, r2); return; }
    float aspect = sqrtf(1.f - 0.9f * anisotropic);
    ax = fmaxf(0.001f, r2 / aspect); ay = fmaxf(0.001f, r2 * aspect);
}

// ---- bitmap textures (src/textures/bitmap.cpp on dr::Texture<Float,2>::eval_nonaccel,
//      ext/drjit/include/drjit/texture.h:484-556,704-735; wrap :1369-1396) -------------------------
LR_DEV int tex_wrap(int p, int r, uint32_t wrap) {
    if (wrap == LR_WRAP_CLAMP) return min(max(p, 0), r - 1);
    int sh = p < 0 ? p + 1 : p; int div = sh / r;
    int mod = p - div * r; if (mod < 0) mod += r;
    if (wrap == LR_WRAP_MIRROR) { bool keep = ((div & 1) == 0) != (p < 0); if (!keep) mod = r - 1 - mod; }
    return mod;
}
LR_DEV V2 tex_xform(const DTex &t, V2 uv) {
    return mk2(fmaf(t.to_uv[0], uv.x, fmaf(t.to_uv[1], uv.y, t.to_uv[2])), fmaf(t.to_uv[3], uv.x, fmaf(t.to_uv[4], uv.y, t.to_uv[5])));
}
LR_DEV float tex_lum(const float *c) { return __fadd_rn(__fadd_rn(__fmul_rn(c[0], 0.212671f), __fmul_rn(c[1], 0.715160f)), __fmul_rn(c[2], 0.072169f)); }
// the four bilinear taps (00, 10, 01, 11) and the fractional weights
LR_DEV void tex_fetch4(const DTex &t, V2 uv, const float *tap[4], float &w1x, float &w1y) {
    float fx = fmaf(uv.x, (float) t.w, -.5f), fy = fmaf(uv.y, (float) t.h, -.5f);
    int ix = (int) floorf(fx), iy = (int) floorf(fy);
    w1x = __fsub_rn(fx, (float) ix); w1y = __fsub_rn(fy, (float) iy);
    int x0 = tex_wrap(ix, t.w, t.wrap), x1 = tex_wrap(ix + 1, t.w, t.wrap), y0 = tex_wrap(iy, t.h, t.wrap), y1 = tex_wrap(iy + 1, t.h, t.wrap);
    tap[0] = t.data + ((size_t) y0 * t.w + x0) * t.ch; tap[1] = t.data + ((size_t) y0 * t.w + x1) * t.ch;
    tap[2] = t.data + ((size_t) y1 * t.w + x0) * t.ch; tap[3] = t.data + ((size_t) y1 * t.w + x1) * t.ch;
}
// Texture::eval for *_rgb variants (bitmap.cpp:443-485): 1 channel => grey
static __device__ __noinline__ V3 tex_eval(const DTex &t, V2 uv_in) {
    V2 uv = tex_xform(t, uv_in); float o[3] = {0.f, 0.f, 0.f};
    if (t.filter == LR_TEX_NEAREST) {
        int ix = tex_wrap((int) floorf(__fmul_rn(uv.x, (float) t.w)), t.w, t.wrap), iy = tex_wrap((int) floorf(__fmul_rn(uv.y, (float) t.h)), t.h, t.wrap);
        const float *p = t.data + ((size_t) iy * t.w + ix) * t.ch;
        for (int c = 0; c < t.ch; ++c) o[c] = __ldg(p + c);
    } else {
        const float *tp[4]; float w1x, w1y; tex_fetch4(t, uv, tp, w1x, w1y);
        float w0x = __fsub_rn(1.f, w1x), w0y = __fsub_rn(1.f, w1y);
        const float wt[4] = {__fmul_rn(w0x, w0y), __fmul_rn(w1x, w0y), __fmul_rn(w0x, w1y), __fmul_rn(w1x, w1y)};
        for (int k = 0; k < 4; ++k) for (int c = 0; c < t.ch; ++c) o[c] = fmaf(__ldg(tp[k] + c), wt[k], o[c]);
    }
    return t.ch == 1 ? mk3(o[0]) : mk3(o[0], o[1], o[2]);
}
// eval_1_grad (bitmap.cpp:509-578)
LR_DEV V2 tex_eval_1_grad(const DTex &t, V2 uv_in) {
    if (t.filter != LR_TEX_BILINEAR) return mk2(0.f, 0.f);
    V2 uv = tex_xform(t, uv_in);
    const float *tp[4]; float w1x, w1y; tex_fetch4(t, uv, tp, w1x, w1y);
    float f[4];
    for (int k = 0; k < 4; ++k) {
        if (t.ch == 1) f[k] = __ldg(tp[k]);
        else { float c[3] = {__ldg(tp[k]), __ldg(tp[k] + 1), __ldg(tp[k] + 2)}; f[k] = tex_lum(c); }
    }
    float w0x = __fsub_rn(1.f, w1x), w0y = __fsub_rn(1.f, w1y);
    float dx = fmaf(w0y, __fsub_rn(f[1], f[0]), __fmul_rn(w1y, __fsub_rn(f[3], f[2]))),
          dy = fmaf(w0x, __fsub_rn(f[2], f[0]), __fmul_rn(w1x, __fsub_rn(f[3], f[1])));
    float du = __fadd_rn(__fmul_rn(t.to_uv[0], dx), __fmul_rn(t.to_uv[3], dy)), dv = __fadd_rn(__fmul_rn(t.to_uv[1], dx), __fmul_rn(t.to_uv[4], dy));
    return mk2(__fmul_rn((float) t.w, du), __fmul_rn((float) t.h, dv));
}

// ---- bumpmap (src/bsdfs/bumpmap.cpp) -------------------------------------------------------------
// The perturbed frame is recomputed where it is needed (before and after the nested BSDF call) from the
// few interaction fields below, so lanes on plain BSDFs carry no extra live state through the stage kernels.
struct BumpIn { V3 n, shn, shs, sht, wi; V2 uv; uint32_t prim; int mesh; };
// BumpMap::frame (bumpmap.cpp:248-276).  dp_du / dp_dv: Mesh::compute_surface_interaction (mesh.cpp:1540-1575),
// recomputed from the triangle with the arithmetic of compute_si.
LR_DEV Frame bump_frame(const DScene &sc, const lr_bsdf &b, const BumpIn in) {
    const DTri tr = sc.tris_gid[in.prim];
    const DMesh &me = sc.meshes[in.mesh];
    V3 p0 = f4xyz(tr.v0), p1 = f4xyz(tr.v1), p2 = f4xyz(tr.v2);
    V3 dp0 = p1 - p0, dp1 = p2 - p0;
    V3 dp_du, dp_dv; coordinate_system(normalize(cross(dp0, dp1)), dp_du, dp_dv);
    if (me.uv) {
        uint32_t face = in.prim - me.tri_offset;
        uint32_t i0 = me.idx[3 * face], i1 = me.idx[3 * face + 1], i2 = me.idx[3 * face + 2];
        V2 uv0 = mk2(me.uv[2 * i0], me.uv[2 * i0 + 1]), uv1 = mk2(me.uv[2 * i1], me.uv[2 * i1 + 1]), uv2 = mk2(me.uv[2 * i2], me.uv[2 * i2 + 1]);
        V2 duv0 = mk2(uv1.x - uv0.x, uv1.y - uv0.y), duv1 = mk2(uv2.x - uv0.x, uv2.y - uv0.y);
        float det = fmaf(duv0.x, duv1.y, -__fmul_rn(duv0.y, duv1.x));
        if (det != 0.f) {
            float inv_det = 1.f / det;
            dp_du = mk3(fmaf(duv1.y, dp0.x, -__fmul_rn(duv0.y, dp1.x)), fmaf(duv1.y, dp0.y, -__fmul_rn(duv0.y, dp1.y)),
                        fmaf(duv1.y, dp0.z, -__fmul_rn(duv0.y, dp1.z))) * inv_det;
            dp_dv = mk3(fmaf(-duv1.x, dp0.x, __fmul_rn(duv0.x, dp1.x)), fmaf(-duv1.x, dp0.y, __fmul_rn(duv0.x, dp1.y)),
                        fmaf(-duv1.x, dp0.z, __fmul_rn(duv0.x, dp1.z))) * inv_det;
        }
    }
    V2 g = tex_eval_1_grad(sc.textures[b.bump_tex - 1], in.uv);
    g.x = __fmul_rn(g.x, b.bump_scale); g.y = __fmul_rn(g.y, b.bump_scale);
    V3 pu = fma3(in.shn, __fsub_rn(g.x, dot(in.shn, dp_du)), dp_du), pv = fma3(in.shn, __fsub_rn(g.y, dot(in.shn, dp_dv)), dp_dv);
    Frame r;
    r.n = normalize(cross(pu, pv));
    if (dot(in.n, r.n) < 0.f) r.n = -r.n;
    r.n = mk3(dot(r.n, in.shs), dot(r.n, in.sht), dot(r.n, in.shn));          // si.to_local
    if (b.bump_flags & LR_BUMP_FLIP_INVALID_NORMALS) {
        if (in.wi.z * dot(in.wi, r.n) <= 0.f) r.n = mk3(-r.n.x, -r.n.y, r.n.z);
    }
    r.s = normalize(fma3(r.n, -dot(r.n, dp_du), dp_du));                       // (sic) world-space dp_du, as the reference
    r.t = cross(r.n, r.s);
    return r;
}
// eval_shadow_terminator (src/bsdfs/normalmap_helpers.h:20-25; Frame::tan_theta_2, frame.h:80-83)
LR_DEV float tan_theta_2(V3 v) { return fmaxf(fmaf(-v.z, v.z, 1.f), 0.f) / sqr(v.z); }
LR_DEV float shadow_terminator(V3 pn, V3 wo) {
    float alpha2 = fminf(0.125f * tan_theta_2(pn), 1.f);
    return 2.f / (1.f + sqrtf(1.f + alpha2 * tan_theta_2(wo)));
}

struct BumpSampleOut { BSDFSample bs; V3 weight; };
static __device__ __noinline__ BumpSampleOut bump_sample(const DScene &sc, const lr_bsdf &b, const BumpIn in, float s1, V2 s2);
static __device__ __noinline__ float4 bump_eval_pdf(const DScene &sc, const lr_bsdf &b, const BumpIn in, V3 wo);

struct PrParams { float anisotropic, roughness, spec_trans, metallic, clearcoat, flatness, sheen, brdf, bsdf; };

struct BSDF {
    const DScene &sc; const lr_bsdf &b; V2 uv; const SurfaceInteraction *si;
    LR_DEV BSDF(const DScene &sc_, const SurfaceInteraction &si_) : sc(sc_), b(sc_.bsdfs[sc_.meshes[si_.mesh].bsdf]), uv(si_.uv), si(&si_) {}
    LR_DEV BSDF(const DScene &sc_, const lr_bsdf &b_, V2 uv_) : sc(sc_), b(b_), uv(uv_), si(nullptr) {}      // nested BSDF of a bump map
    LR_DEV BumpIn bump_in() const {
        BumpIn in; in.n = si->n; in.shn = si->sh.n; in.shs = si->sh.s; in.sht = si->sh.t; in.wi = si->wi; in.uv = si->uv; in.prim = si->prim; in.mesh = si->mesh;
        return in;
    }
    // Bump-mapped BSDFs leave the stage kernel through one out-of-line call, so lanes on plain BSDFs keep the
    // register allocation they had (the wrapper's state never lives across the inlined nested code).
    LR_DEV BSDFSample sample(V3 wi, float s1, V2 s2, V3 &weight) const {
        if (b.bump_tex) { BumpSampleOut o = bump_sample(sc, b, bump_in(), s1, s2); weight = o.weight; return o.bs; }
        return sample_n(wi, s1, s2, weight);
    }
    LR_DEV V3 eval_pdf(V3 wi, V3 wo, float &pdf) const {
        if (b.bump_tex) { float4 r = bump_eval_pdf(sc, b, bump_in(), wo); pdf = r.w; return mk3(r.x, r.y, r.z); }
        return eval_pdf_n(wi, wo, pdf);
    }
    LR_DEV V3 refl() const {
        if (b.tex_kind == 2) return tex_eval(sc.textures[b.refl_tex - 1], uv);   // bitmap reflectance
        if (b.tex_kind == 1) {                      // checkerboard (src/textures/checkerboard.cpp:70-89)
            float u = fmaf(b.tex_to_uv[0], uv.x, fmaf(b.tex_to_uv[1], uv.y, b.tex_to_uv[2])),
                  v = fmaf(b.tex_to_uv[3], uv.x, fmaf(b.tex_to_uv[4], uv.y, b.tex_to_uv[5]));
            bool mx = u - floorf(u) > .5f, my = v - floorf(v) > .5f;
            if (mx != my) return mk3(b.tex_color1[0], b.tex_color1[1], b.tex_color1[2]);
        }
        return mk3(b.reflectance[0], b.reflectance[1], b.reflectance[2]);
    }
    LR_DEV V3 trans() const { return mk3(b.transmittance[0], b.transmittance[1], b.transmittance[2]); }
    LR_DEV bool has(uint32_t f) const { return (b.flags & f) != 0; }
    LR_DEV uint32_t flags() const {
        switch (b.kind) {
            case LR_BSDF_DIFFUSE: return F_DIFFUSE_REFL;
            case LR_BSDF_ROUGHDIELECTRIC: return F_GLOSSY_REFL | F_GLOSSY_TRANS;
            case LR_BSDF_PRINCIPLED: return F_DIFFUSE_REFL | F_GLOSSY_REFL | (has(LR_PR_HAS_SPEC_TRANS) ? F_GLOSSY_TRANS : 0u);
            case LR_BSDF_NULL: return F_NULL;
            case LR_BSDF_DIELECTRIC: return F_DELTA_REFL | F_DELTA_TRANS;
        }
        return 0;
    }
    LR_DEV float eval_null_transmission() const { return b.kind == LR_BSDF_NULL ? 1.f : 0.f; }

    LR_DEV PrParams pr() const {
        PrParams p;
        p.anisotropic = has(LR_PR_HAS_ANISOTROPIC) ? b.anisotropic : 0.f; p.roughness = b.roughness;
        p.spec_trans = has(LR_PR_HAS_SPEC_TRANS) ? b.spec_trans : 0.f; p.metallic = has(LR_PR_HAS_METALLIC) ? b.metallic : 0.f;
        p.clearcoat = has(LR_PR_HAS_CLEARCOAT) ? b.clearcoat : 0.f; p.flatness = has(LR_PR_HAS_FLATNESS) ? b.flatness : 0.f;
        p.sheen = has(LR_PR_HAS_SHEEN) ? b.sheen : 0.f;
        p.brdf = (1.f - p.metallic) * (1.f - p.spec_trans); p.bsdf = (1.f - p.metallic) * p.spec_trans;
        return p;
    }
    LR_DEV void pr_probs(const PrParams &p, bool front, float F, float &ps_r, float &ps_t, float &p_cc, float &p_d) const {
        ps_r = front ? b.main_spec_srate * (1.f - p.bsdf * (1.f - F)) : F;
        ps_t = has(LR_PR_HAS_SPEC_TRANS) ? (front ? b.main_spec_srate * p.bsdf * (1.f - F) : (1.f - F)) : 0.f;
        p_cc = has(LR_PR_HAS_CLEARCOAT) ? (front ? 0.25f * p.clearcoat * b.clearcoat_srate : 0.f) : 0.f;
        p_d = front ? p.brdf * b.diff_refl_srate : 0.f;
    }
    LR_DEV V3 principled_fresnel(float F_diel, float metallic, float spec_tint, V3 base, float lum, float cos_theta_i, bool front, float bsdf_) const {
        bool outside = cos_theta_i >= 0.f; float rcp_eta = 1.f / b.eta, eta_it = outside ? b.eta : rcp_eta;
        V3 F_schlick = mk3(0.f);
        if (has(LR_PR_HAS_METALLIC)) F_schlick += metallic * calc_schlick3(base, cos_theta_i, b.eta);
        if (has(LR_PR_HAS_SPEC_TINT)) {
            V3 c_tint = lum > 0.f ? base / lum : mk3(1.f);
            V3 F0 = c_tint * sqr((eta_it - 1.f) / (eta_it + 1.f));
            F_schlick += (1.f - metallic) * spec_tint * calc_schlick3(F0, cos_theta_i, b.eta);
        }
        V3 F_front = mk3((1.f - metallic) * (1.f - spec_tint) * F_diel) + F_schlick;
        return front ? F_front : mk3(bsdf_ * F_diel);
    }

    LR_DEV V3 eval_principled(V3 wi, V3 wo) const {
        float cos_theta_i = wi.z; if (cos_theta_i == 0.f) return mk3(0.f);
        PrParams p = pr(); V3 base = refl(); float cos_theta_o = wo.z;
        bool refl_ = cos_theta_i * cos_theta_o > 0.f, refr = cos_theta_i * cos_theta_o < 0.f, front = cos_theta_i > 0.f;
        float inv_eta = 1.f / b.eta, eta_path = front ? b.eta : inv_eta, inv_eta_path = front ? inv_eta : b.eta;
        float ax, ay; calc_dist_params(p.anisotropic, p.roughness, has(LR_PR_HAS_ANISOTROPIC), ax, ay);
        Microfacet sd; sd.init(LR_MF_GGX, ax, ay, true);
        V3 wh = normalize(wi + wo * (refl_ ? 1.f : eta_path)); wh = mulsign(wh, wh.z);
        float F, ct, eit, eti; fresnel(dot(wi, wh), b.eta, F, ct, eit, eti);
        bool rc = mac_mic_compat(wh, wi, wo, cos_theta_i, true), tc = mac_mic_compat(wh, wi, wo, cos_theta_i, false);
        bool a_sr = refl_ && rc && F > 0.f;
        bool a_cc = has(LR_PR_HAS_CLEARCOAT) && p.clearcoat > 0.f && refl_ && rc && front;
        bool a_st = has(LR_PR_HAS_SPEC_TRANS) && p.bsdf > 0.f && refr && tc && F < 1.f;
        bool a_d = p.brdf > 0.f && refl_ && front;
        bool a_sh = has(LR_PR_HAS_SHEEN) && p.sheen > 0.f && refl_ && (1.f - p.metallic > 0.f) && front;
        float D = sd.eval(wh), G = sd.G(wi, wo, wh);
        V3 value = mk3(0.f);
        if (a_sr) {
            float lum = has(LR_PR_HAS_SPEC_TINT) ? luminance(base) : 1.f, spec_tint = has(LR_PR_HAS_SPEC_TINT) ? b.spec_tint : 0.f;
            V3 Fp = principled_fresnel(F, p.metallic, spec_tint, base, lum, dot(wi, wh), front, p.bsdf);
            value += Fp * D * G / (4.f * fabsf(cos_theta_i));
        }
        if (a_st) {
            float scale = sqr(inv_eta_path);
            value += mk3(sqrtf(base.x), sqrtf(base.y), sqrtf(base.z)) * p.bsdf *
                     fabsf((scale * (1.f - F) * D * G * eta_path * eta_path * dot(wi, wh) * dot(wo, wh)) /
                           (cos_theta_i * sqr(dot(wi, wh) + eta_path * dot(wo, wh))));
        }
        if (a_cc) {
            float Fcc = calc_schlick1(0.04f, dot(wi, wh), b.eta), al = lerpf(0.1f, 0.001f, b.clearcoat_gloss);
            float Dcc = gtr1_eval(al, wh), Gcc = smith_ggx1(wi, wh, 0.25f) * smith_ggx1(wo, wh, 0.25f);
            value += mk3((p.clearcoat * 0.25f) * Fcc * Dcc * Gcc * fabsf(cos_theta_o));
        }
        if (a_d) {
            float Fo = schlick_weight(fabsf(cos_theta_o)), Fi = schlick_weight(fabsf(cos_theta_i));
            float f_diff = (1.f - 0.5f * Fi) * (1.f - 0.5f * Fo);
            float cos_theta_d = dot(wh, wo), Rr = 2.f * p.roughness * sqr(cos_theta_d);
            float f_retro = Rr * (Fo + Fi + Fo * Fi * (Rr - 1.f));
            if (has(LR_PR_HAS_FLATNESS)) {
                float Fss90 = Rr / 2.f, Fss = lerpf(1.f, Fss90, Fo) * lerpf(1.f, Fss90, Fi);
                float f_ss = 1.25f * (Fss * (1.f / (fabsf(cos_theta_o) + fabsf(cos_theta_i)) - 0.5f) + 0.5f);
                value += p.brdf * fabsf(cos_theta_o) * base * kInvPi * lerpf(f_diff + f_retro, f_ss, p.flatness);
            } else value += p.brdf * fabsf(cos_theta_o) * base * kInvPi * (f_diff + f_retro);
            if (a_sh) {
                float Fd = schlick_weight(fabsf(cos_theta_d));
                if (has(LR_PR_HAS_SHEEN_TINT)) {
                    float lum = luminance(base); V3 c_tint = lum > 0.f ? base / lum : mk3(1.f);
                    V3 c_sheen = mk3(lerpf(1.f, c_tint.x, b.sheen_tint), lerpf(1.f, c_tint.y, b.sheen_tint), lerpf(1.f, c_tint.z, b.sheen_tint));
                    value += p.sheen * (1.f - p.metallic) * Fd * c_sheen * fabsf(cos_theta_o);
                } else value += mk3(p.sheen * (1.f - p.metallic) * Fd * fabsf(cos_theta_o));
            }
        }
        return value;
    }
    LR_DEV float pdf_principled(V3 wi, V3 wo) const {
        float cos_theta_i = wi.z; if (cos_theta_i == 0.f) return 0.f;
        PrParams p = pr(); bool front = cos_theta_i > 0.f;
        float eta_path = front ? b.eta : 1.f / b.eta, cos_theta_o = wo.z;
        bool refl_ = cos_theta_i * cos_theta_o > 0.f, refr = cos_theta_i * cos_theta_o < 0.f;
        V3 wh = normalize(wi + wo * (refl_ ? 1.f : eta_path)); wh = mulsign(wh, wh.z);
        float ax, ay; calc_dist_params(p.anisotropic, p.roughness, has(LR_PR_HAS_ANISOTROPIC), ax, ay);
        Microfacet sd; sd.init(LR_MF_GGX, ax, ay, true);
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
        float spdf = sd.pdf(mulsign(wi, cos_theta_i), wh);
        if (mr) pdf += ps_r * spdf * dwh_dwo_abs;
        if (refl_) pdf += p_d * (kInvPi * wo.z);
        if (has(LR_PR_HAS_SPEC_TRANS)) {
            bool mt = mac_mic_compat(wh, wi, wo, cos_theta_i, false) && refr;
            if (mt) pdf += ps_t * spdf * dwh_dwo_abs;
        }
        if (has(LR_PR_HAS_CLEARCOAT) && mr) pdf += p_cc * gtr1_pdf(lerpf(0.1f, 0.001f, b.clearcoat_gloss), wh) * dwh_dwo_abs;
        return pdf;
    }
    LR_DEV BSDFSample sample_principled(V3 wi, float s1, V2 s2, V3 &weight) const {
        BSDFSample bs; bs.wo = mk3(0.f); bs.pdf = 0.f; bs.eta = 0.f; bs.sampled_type = 0; weight = mk3(0.f);
        float cos_theta_i = wi.z; bool active = cos_theta_i != 0.f;
        if (!active) return bs;
        PrParams p = pr(); bool front = cos_theta_i > 0.f;
        float ax, ay; calc_dist_params(p.anisotropic, p.roughness, has(LR_PR_HAS_ANISOTROPIC), ax, ay);
        Microfacet sd; sd.init(LR_MF_GGX, ax, ay, true);
        V3 m_spec; float pdf_m; sd.sample(mulsign(wi, cos_theta_i), s2, m_spec, pdf_m);
        float F, cos_theta_t, eta_it, eta_ti; fresnel(dot(wi, m_spec), b.eta, F, cos_theta_t, eta_it, eta_ti);
        active &= (front || p.bsdf > 0.f);
        float ps_r, ps_t, p_cc, p_d; pr_probs(p, front, F, ps_r, ps_t, p_cc, p_d);
        float rcp_tot = 1.f / (ps_r + ps_t + p_cc + p_d);
        ps_t *= rcp_tot; p_cc *= rcp_tot; p_d *= rcp_tot;
        float curr = 0.f;
        bool s_diff = active && s1 < p_d; curr += p_d;
        bool s_cc = has(LR_PR_HAS_CLEARCOAT) && active && s1 >= curr && s1 < curr + p_cc; curr += p_cc;
        bool s_tr = has(LR_PR_HAS_SPEC_TRANS) && active && s1 >= curr && s1 < curr + ps_t; curr += ps_t;
        bool s_rf = active && s1 >= curr;
        bs.eta = 1.f;
        if (s_rf) { V3 wo = reflect_m(wi, m_spec); bs.wo = wo; bs.sampled_type = F_GLOSSY_REFL;
            active &= mac_mic_compat(m_spec, wi, wo, cos_theta_i, true) && (cos_theta_i * wo.z > 0.f); }
        if (s_tr) { V3 wo = refract_m(wi, m_spec, cos_theta_t, eta_ti); bs.wo = wo; bs.sampled_type = F_GLOSSY_TRANS; bs.eta = eta_it;
            active &= mac_mic_compat(m_spec, wi, wo, cos_theta_i, false) && (cos_theta_i * wo.z < 0.f); }
        if (s_cc) { V3 m_cc = gtr1_sample(lerpf(0.1f, 0.001f, b.clearcoat_gloss), s2); V3 wo = reflect_m(wi, m_cc); bs.wo = wo; bs.sampled_type = F_GLOSSY_REFL;
            active &= mac_mic_compat(m_cc, wi, wo, cos_theta_i, true) && (cos_theta_i * wo.z > 0.f); }
        if (s_diff) { V3 wo = square_to_cosine_hemisphere(s2); bs.wo = wo; bs.sampled_type = F_DIFFUSE_REFL; active &= cos_theta_i * wo.z > 0.f; }
        bs.pdf = active ? pdf_principled(wi, bs.wo) : 0.f;
        active &= bs.pdf > 0.f;
        if (active) weight = eval_principled(wi, bs.wo) / bs.pdf;
        return bs;
    }

    LR_DEV BSDFSample sample_n(V3 wi, float s1, V2 s2, V3 &weight) const {
        BSDFSample bs; bs.wo = mk3(0.f); bs.pdf = 0.f; bs.eta = 0.f; bs.sampled_type = 0; weight = mk3(0.f);
        switch (b.kind) {
        case LR_BSDF_DIFFUSE: {
            if (!(wi.z > 0.f)) return bs;
            bs.wo = square_to_cosine_hemisphere(s2); bs.pdf = kInvPi * bs.wo.z; bs.eta = 1.f; bs.sampled_type = F_DIFFUSE_REFL;
            if (bs.pdf > 0.f) weight = refl();
            return bs;
        }
        case LR_BSDF_NULL: { bs.wo = -wi; bs.sampled_type = F_NULL; bs.eta = 1.f; bs.pdf = 1.f; weight = mk3(1.f); return bs; }
        case LR_BSDF_DIELECTRIC: {
            float r_i, ct, eta_it, eta_ti; fresnel(wi.z, b.eta, r_i, ct, eta_it, eta_ti);
            float t_i = 1.f - r_i; bool sel_r = s1 <= r_i;
            bs.pdf = sel_r ? r_i : t_i; bs.sampled_type = sel_r ? F_DELTA_REFL : F_DELTA_TRANS;
            bs.wo = sel_r ? reflect_l(wi) : refract_l(wi, ct, eta_ti); bs.eta = sel_r ? 1.f : eta_it;
            weight = sel_r ? refl() : trans() * sqr(eta_ti);
            return bs;
        }
        case LR_BSDF_ROUGHDIELECTRIC: {
            float cos_theta_i = wi.z; bool active = cos_theta_i != 0.f;
            Microfacet distr; distr.init(b.mf_kind, b.alpha_u, b.alpha_v, b.sample_visible != 0); Microfacet sd = distr;
            if (!b.sample_visible) sd.scale_alpha(1.2f - .2f * sqrtf(fabsf(cos_theta_i)));
            V3 m; sd.sample(mulsign(wi, cos_theta_i), s2, m, bs.pdf);
            active &= bs.pdf != 0.f;
            float F, cos_theta_t, eta_it, eta_ti; fresnel(dot(wi, m), b.eta, F, cos_theta_t, eta_it, eta_ti);
            bool sel_r = s1 <= F && active, sel_t = !sel_r && active;
            V3 w = mk3(1.f);
            bs.pdf *= sel_r ? F : 1.f - F;
            bs.eta = sel_r ? 1.f : eta_it; bs.sampled_type = sel_r ? F_GLOSSY_REFL : F_GLOSSY_TRANS;
            float dwh_dwo = 0.f;
            if (sel_r) { bs.wo = reflect_m(wi, m); w *= refl(); dwh_dwo = 1.f / (4.f * dot(bs.wo, m)); }
            if (sel_t) {
                bs.wo = refract_m(wi, m, cos_theta_t, eta_ti); w *= trans() * sqr(eta_ti);
                dwh_dwo = (sqr(bs.eta) * dot(bs.wo, m)) / sqr(dot(wi, m) + bs.eta * dot(bs.wo, m));
            }
            if (b.sample_visible) w *= distr.smith_g1(bs.wo, m);
            else w *= distr.G(wi, bs.wo, m) * dot(wi, m) / (cos_theta_i * m.z);
            bs.pdf *= fabsf(dwh_dwo);
            weight = active ? w : mk3(0.f);
            return bs;
        }
        case LR_BSDF_PRINCIPLED: return sample_principled(wi, s1, s2, weight);
        }
        return bs;
    }
    // eval and pdf in one go (the integrators always need both)
    LR_DEV V3 eval_pdf_n(V3 wi, V3 wo, float &pdf) const {
        pdf = 0.f;
        switch (b.kind) {
        case LR_BSDF_DIFFUSE: {
            if (!(wi.z > 0.f && wo.z > 0.f)) return mk3(0.f);
            pdf = kInvPi * wo.z; return refl() * kInvPi * wo.z;
        }
        case LR_BSDF_ROUGHDIELECTRIC: {
            float ci = wi.z, co = wo.z; if (ci == 0.f) return mk3(0.f);
            bool refl_ = ci * co > 0.f;
            float eta = ci > 0.f ? b.eta : 1.f / b.eta, inv_eta = ci > 0.f ? 1.f / b.eta : b.eta;
            V3 m = normalize(wi + wo * (refl_ ? 1.f : eta)); m = mulsign(m, m.z);
            Microfacet distr; distr.init(b.mf_kind, b.alpha_u, b.alpha_v, b.sample_visible != 0);
            float D = distr.eval(m), F, ct, eit, eti; fresnel(dot(wi, m), b.eta, F, ct, eit, eti);
            float G = distr.G(wi, wo, m);
            // pdf (roughdielectric.cpp:436-499)
            bool pa = dot(wi, m) * wi.z > 0.f && dot(wo, m) * wo.z > 0.f;
            float dwh_dwo = refl_ ? 1.f / (4.f * dot(wo, m)) : (eta * eta * dot(wo, m)) / sqr(dot(wi, m) + eta * dot(wo, m));
            Microfacet sd = distr; if (!b.sample_visible) sd.scale_alpha(1.2f - .2f * sqrtf(fabsf(wi.z)));
            float prob = sd.pdf(mulsign(wi, wi.z), m) * (refl_ ? F : 1.f - F);
            pdf = pa ? prob * fabsf(dwh_dwo) : 0.f;
            if (refl_) return refl() * (F * D * G / (4.f * fabsf(ci)));
            float scale = sqr(inv_eta);
            float v = fabsf((scale * (1.f - F) * D * G * eta * eta * dot(wi, m) * dot(wo, m)) / (ci * sqr(dot(wi, m) + eta * dot(wo, m))));
            return trans() * v;
        }
        case LR_BSDF_PRINCIPLED: pdf = pdf_principled(wi, wo); return eval_principled(wi, wo);
        default: return mk3(0.f);
        }
    }
};

// BumpMap::sample / eval / pdf (src/bsdfs/bumpmap.cpp:136-233): the nested BSDF runs in the perturbed frame; `in.wi`
// is the incident direction in the original shading frame.
static __device__ __noinline__ BumpSampleOut bump_sample(const DScene &sc, const lr_bsdf &b, const BumpIn in, float s1, V2 s2) {
    Frame pf = bump_frame(sc, b, in); BSDF nested(sc, b, in.uv); BumpSampleOut o;
    o.bs = nested.sample_n(pf.to_local(in.wi), s1, s2, o.weight);
    bool active = o.weight.x != 0.f || o.weight.y != 0.f || o.weight.z != 0.f;
    V3 pwo = pf.to_world(o.bs.wo);
    active &= o.bs.wo.z * pwo.z > 0.f;
    o.bs.wo = pwo;
    if (b.bump_flags & LR_BUMP_SHADOWING_FUNCTION) o.weight = o.weight * shadow_terminator(pf.n, pwo);
    if (!active) o.weight = mk3(0.f);
    return o;
}
static __device__ __noinline__ float4 bump_eval_pdf(const DScene &sc, const lr_bsdf &b, const BumpIn in, V3 wo) {
    Frame pf = bump_frame(sc, b, in); V3 pwo = pf.to_local(wo);
    if (!(wo.z * pwo.z > 0.f)) return make_float4(0.f, 0.f, 0.f, 0.f);
    BSDF nested(sc, b, in.uv); float pdf;
    V3 v = nested.eval_pdf_n(pf.to_local(in.wi), pwo, pdf);
    if (b.bump_flags & LR_BUMP_SHADOWING_FUNCTION) v = v * shadow_terminator(pf.n, wo);
    return make_float4(v.x, v.y, v.z, pdf);
}

// ---- phase functions -----------------------------------------------------------------------
struct Phase {
    uint32_t kind; float g;
    LR_DEV float eval_hg(float cos_theta) const {
        float temp = 1.f + sqr(g) + 2.f * g * cos_theta;
        return kInvFourPi * (1.f - sqr(g)) / (temp * sqrtf(temp));
    }
    LR_DEV V3 sample(V3 wi, V2 u, float &pdf) const {
        if (kind == LR_PHASE_ISOTROPIC) { pdf = kInvFourPi; return square_to_uniform_sphere(u); }
        float sqr_term = (1.f - sqr(g)) / (1.f - g + 2.f * g * u.x), cos_theta = (1.f + sqr(g) - sqr(sqr_term)) / (2.f * g);
        if (fabsf(g) < kEps) cos_theta = 1.f - 2.f * u.x;
        float sin_theta = safe_sqrt(1.f - sqr(cos_theta)), sp, cp; sincosf(2.f * kPi * u.y, &sp, &cp);
        Frame f = frame_from_n(wi);
        pdf = eval_hg(-cos_theta);
        return f.to_world(mk3(sin_theta * cp, sin_theta * sp, -cos_theta));
    }
    LR_DEV float eval_pdf(V3 wi, V3 wo) const { return kind == LR_PHASE_ISOTROPIC ? kInvFourPi : eval_hg(dot(wo, wi)); }
};

} // namespace lr
