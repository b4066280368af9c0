// lr_emitter.cuh -- emitters used for NEE + MIS on the hot path (device code).
// area (src/emitters/area.cpp:83-197 + src/render/shape.cpp:344-374 + src/shapes/rectangle.cpp:181-199),
// constant (src/emitters/constant.cpp:92-163), envmap (src/emitters/envmap.cpp:353-476, 527-580) with the
// hierarchical 2D warp (include/mitsuba/core/distr_2d.h:371-830), scene-level selection
// (src/render/scene.cpp:265-406).
#pragma once
#include "lr_scene.cuh"

namespace lr {

struct DirectionSample { V3 p, n, d; V2 uv; float pdf, dist; bool delta; int emitter; };

LR_DEV float clip01(float v) { return fmaxf(fminf(v, 1.f), 0.f); }
LR_DEV float interval_to_linear(float v0, float v1, float s) {          // warp.h:446-453
    return fabsf(v0 - v1) > 1e-4f * (v0 + v1) ? (v0 - safe_sqrt(lerpf(sqr(v0), sqr(v1), s))) / (v0 - v1) : s;
}
LR_DEV uint32_t h2d_index(uint32_t x, uint32_t y, uint32_t width) {   // distr_2d.h Level::index: 2x2 patches contiguous
    return ((x & 1u) | (((x & ~1u) | (y & 1u)) << 1)) + ((y & ~1u) * width);
}
// Hierarchical2D::sample
LR_DEV V2 h2d_sample(const DEmitter &em, V2 s, float &pdf) {
    s.x = clip01(s.x); s.y = clip01(s.y);
    uint32_t ox = 0, oy = 0;
    for (int l = em.n_levels - 2; l > 0; --l) {
        const float *lv = em.warp_data + em.level_off[l];
        ox <<= 1; oy <<= 1;
        const float4 v = *reinterpret_cast<const float4 *>(lv + h2d_index(ox, oy, (uint32_t) em.level_w[l]));   // v00 v10 v01 v11
        s.x = clip01(s.x); s.y = clip01(s.y);
        float r0 = v.x + v.y, r1 = v.z + v.w;
        s.y *= r0 + r1;
        bool mask = s.y > r0;
        if (mask) { oy += 1u; s.y -= r0; }
        s.y /= mask ? r1 : r0;
        float c0 = mask ? v.z : v.x, c1 = mask ? v.w : v.y;
        s.x *= c0 + c1;
        mask = s.x > c0;
        if (mask) { s.x -= c0; ox += 1u; }
        s.x /= mask ? c1 : c0;
    }
    const float *l0 = em.warp_data + em.level_off[0]; uint32_t w0 = (uint32_t) em.level_w[0];
    uint32_t oi = ox + oy * w0;
    float v00 = l0[oi], v10 = l0[oi + 1], v01 = l0[oi + w0], v11 = l0[oi + w0 + 1];
    float r0 = v00 + v10, r1 = v01 + v11;                                   // warp::square_to_bilinear
    s.y = interval_to_linear(r0, r1, s.y);
    float c0 = lerpf(v00, v01, s.y), c1 = lerpf(v10, v11, s.y);
    s.x = interval_to_linear(c0, c1, s.x);
    pdf = lerpf(c0, c1, s.x);
    float psx = 1.f / (float) (w0 - 1), psy = 1.f / (float) (em.e.height - 1);
    return mk2(((float) (int) ox + s.x) * psx, ((float) (int) oy + s.y) * psy);
}
LR_DEV float h2d_eval(const DEmitter &em, V2 pos) {
    uint32_t w0 = (uint32_t) em.level_w[0], npx = w0 - 1, npy = em.e.height - 1;
    pos.x = clip01(pos.x) * (float) npx; pos.y = clip01(pos.y) * (float) npy;
    uint32_t ox = min((uint32_t) (int) pos.x, npx - 1), oy = min((uint32_t) (int) pos.y, npy - 1);
    pos.x -= (float) (int) ox; pos.y -= (float) (int) oy;
    const float *l0 = em.warp_data + em.level_off[0]; uint32_t oi = ox + oy * w0;
    float v00 = l0[oi], v10 = l0[oi + 1], v01 = l0[oi + w0], v11 = l0[oi + w0 + 1];
    return lerpf(lerpf(v00, v10, pos.x), lerpf(v01, v11, pos.x), pos.y);
}
LR_DEV float safe_acos(float x) { return acosf(fminf(fmaxf(x, -1.f), 1.f)); }

LR_DEV V3 env_eval_uv(const DEmitter &em, V2 uv) {                          // envmap.cpp:527-580 (rgb branch)
    uint32_t rx = em.e.width + 1, ry = em.e.height;
    uv.x -= .5f / (float) (rx - 1u);
    uv.x -= floorf(uv.x); uv.y -= floorf(uv.y);
    uv.x *= (float) (rx - 1u); uv.y *= (float) (ry - 1u);
    uint32_t px = min((uint32_t) uv.x, rx - 2u), py = min((uint32_t) uv.y, ry - 2u);
    float w1x = uv.x - (float) px, w1y = uv.y - (float) py, w0x = 1.f - w1x, w0y = 1.f - w1y;
    size_t index = (size_t) py * rx + px; const float *d = em.pixels;
    V3 r;
    float *rp = &r.x;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        float v00 = d[index * 3 + c], v10 = d[(index + 1) * 3 + c], v01 = d[(index + rx) * 3 + c], v11 = d[(index + rx + 1) * 3 + c];
        float v0 = fmaf(w0x, v00, __fmul_rn(w1x, v10)), v1 = fmaf(w0x, v01, __fmul_rn(w1x, v11));
        rp[c] = fmaf(w0y, v0, __fmul_rn(w1y, v1)) * em.e.scale;
    }
    return r;
}

LR_DEV float env_bs_radius(const DScene &sc) { return fmaxf(kRayEps, sc.bs_radius * (1.f + kRayEps)); }

// Emitter::eval(si).  `wi_local_z`: cos_theta(si.wi); `d_world`: direction of travel (= -si.wi for escaped rays)
LR_DEV V3 emitter_eval(const DScene &sc, int id, float wi_local_z, V3 d_world) {
    const DEmitter &em = sc.emitters[id];
    switch (em.e.kind) {
        case LR_EMITTER_AREA: return wi_local_z > 0.f ? mk3(em.e.radiance[0], em.e.radiance[1], em.e.radiance[2]) : mk3(0.f);
        case LR_EMITTER_CONSTANT: return mk3(em.e.radiance[0], em.e.radiance[1], em.e.radiance[2]);
        default: {
            V3 v = xform_vector(em.to_local, d_world);
            return env_eval_uv(em, mk2(atan2f(v.x, -v.z) * kInvTwoPi, safe_acos(v.y) * kInvPi));
        }
    }
}

LR_DEV DirectionSample emitter_sample_direction(const DScene &sc, int id, V3 ref_p, V2 u, V3 &spec) {
    const DEmitter &em = sc.emitters[id]; const lr_emitter &e = em.e;
    DirectionSample ds; ds.emitter = id; ds.delta = false; ds.pdf = 0.f; ds.dist = 0.f; ds.uv = u; spec = mk3(0.f);
    ds.p = ds.n = ds.d = mk3(0.f);
    switch (e.kind) {
    case LR_EMITTER_AREA: {
        if (e.is_rectangle) {
            float a = fmaf(u.x, 2.f, -1.f), b = fmaf(u.y, 2.f, -1.f);
            V3 o = mk3(e.rect_origin[0], e.rect_origin[1], e.rect_origin[2]), eu = mk3(e.rect_edge_u[0], e.rect_edge_u[1], e.rect_edge_u[2]),
               ev = mk3(e.rect_edge_v[0], e.rect_edge_v[1], e.rect_edge_v[2]);
            ds.p = fma3(eu, a, fma3(ev, b, o)); ds.n = mk3(e.rect_normal[0], e.rect_normal[1], e.rect_normal[2]); ds.pdf = e.inv_area;
        } else {
            const DMesh &me = sc.meshes[e.mesh];
            float target = u.y * me.area; uint32_t lo = 0, hi = me.n_faces;           // upper_bound on the area cdf
            while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if (me.area_cdf[mid] <= target) lo = mid + 1; else hi = mid; }
            uint32_t f = min(lo, me.n_faces - 1);
            float c0 = f ? me.area_cdf[f - 1] : 0.f, c1 = me.area_cdf[f];
            V2 b = square_to_uniform_triangle(mk2(u.x, (target - c0) / (c1 - c0)));
            uint32_t i0 = me.idx[3 * f], i1 = me.idx[3 * f + 1], i2 = me.idx[3 * f + 2];
            V3 p0 = mk3(me.pos[3 * i0], me.pos[3 * i0 + 1], me.pos[3 * i0 + 2]), p1 = mk3(me.pos[3 * i1], me.pos[3 * i1 + 1], me.pos[3 * i1 + 2]),
               p2 = mk3(me.pos[3 * i2], me.pos[3 * i2 + 1], me.pos[3 * i2 + 2]);
            V3 e0 = p1 - p0, e1 = p2 - p0;
            ds.p = fma3(e0, b.x, fma3(e1, b.y, p0)); ds.n = normalize(cross(e0, e1)); ds.pdf = 1.f / me.area; ds.uv = b;
        }
        ds.d = ds.p - ref_p; float dist2 = sqnorm(ds.d); ds.dist = sqrtf(dist2); ds.d = ds.d / ds.dist;
        float dp = abs_dot(ds.d, ds.n), x = dist2 / dp;
        ds.pdf *= isfinite(x) ? x : 0.f;
        if (dot(ds.d, ds.n) < 0.f && ds.pdf != 0.f) spec = mk3(e.radiance[0], e.radiance[1], e.radiance[2]) / ds.pdf;
        return ds;
    }
    case LR_EMITTER_CONSTANT: {
        V3 d = square_to_uniform_sphere(u);
        float dist = 2.f * fmaxf(env_bs_radius(sc), norm(ref_p - sc.bs_center));
        ds.p = fma3(d, dist, ref_p); ds.n = -d; ds.pdf = kInvFourPi; ds.d = d; ds.dist = dist;
        spec = mk3(e.radiance[0], e.radiance[1], e.radiance[2]) / ds.pdf;
        return ds;
    }
    default: {
        float pdf; V2 uv = h2d_sample(em, u, pdf);
        uv.x += .5f / (float) (e.width);
        bool active = pdf > 0.f;
        float st, ct, sp, cp; sincosf(uv.y * kPi, &st, &ct); sincosf(uv.x * (2.f * kPi), &sp, &cp);
        V3 d = mk3(sp * st, ct, -(cp * st));
        float dist = 2.f * fmaxf(env_bs_radius(sc), norm(ref_p - sc.bs_center));
        float inv_sin_theta = 1.f / sqrtf(fmaxf(sqr(d.x) + sqr(d.z), sqr(kEps)));
        d = xform_vector(em.to_world, d);
        ds.p = ref_p + d * dist; ds.n = -d; ds.uv = uv;
        ds.pdf = active ? pdf * inv_sin_theta * (1.f / (2.f * sqr(kPi))) : 0.f;
        ds.d = d; ds.dist = dist;
        if (active) spec = env_eval_uv(em, uv) / ds.pdf;
        return ds;
    }
    }
}

LR_DEV float emitter_pdf_direction(const DScene &sc, int id, const DirectionSample &ds) {
    const DEmitter &em = sc.emitters[id]; const lr_emitter &e = em.e;
    switch (e.kind) {
    case LR_EMITTER_AREA: {
        float dp = dot(ds.d, ds.n); if (!(dp < 0.f)) return 0.f;
        float pdf = e.is_rectangle ? e.inv_area : 1.f / sc.meshes[e.mesh].area, adp = fabsf(dp);
        return pdf * (adp != 0.f ? (ds.dist * ds.dist) / adp : 0.f);
    }
    case LR_EMITTER_CONSTANT: return kInvFourPi;
    default: {
        V3 d = xform_vector(em.to_local, ds.d);
        V2 uv = mk2(atan2f(d.x, -d.z) * kInvTwoPi, safe_acos(d.y) * kInvPi);
        uv.x -= .5f / (float) (e.width);
        uv.x -= floorf(uv.x); uv.y -= floorf(uv.y);
        float inv_sin_theta = 1.f / sqrtf(fmaxf(sqr(d.x) + sqr(d.z), sqr(kEps)));
        return h2d_eval(em, uv) * inv_sin_theta * (1.f / (2.f * sqr(kPi)));
    }
    }
}

// Scene::sample_emitter_direction without the visibility test (scene.cpp:333-384, JIT branch)
LR_DEV DirectionSample sample_emitter_direction(const DScene &sc, V3 ref_p, V2 sample, V3 &spec) {
    uint32_t n = sc.n_emitters; DirectionSample ds; spec = mk3(0.f);
    if (n == 0) { ds.pdf = 0.f; ds.emitter = -1; ds.delta = false; ds.d = ds.p = ds.n = mk3(0.f); ds.dist = 0.f; ds.uv = sample; return ds; }
    uint32_t index = 0; float weight = 1.f;
    if (n > 1) {
        float scaled = sample.x * (float) n;
        index = min((uint32_t) scaled, n - 1u); weight = (float) n; sample.x = scaled - (float) index;
    }
    ds = emitter_sample_direction(sc, (int) index, ref_p, sample, spec);
    ds.pdf *= 1.f / (float) n; spec *= weight;
    return ds;
}
LR_DEV float pdf_emitter_direction(const DScene &sc, const DirectionSample &ds) {   // scene.cpp:395-406
    if (ds.emitter < 0) return 0.f;
    return emitter_pdf_direction(sc, ds.emitter, ds) * (1.f / (float) sc.n_emitters);
}
// DirectionSample(scene, si, ref) (include/mitsuba/render/records.h:173-180)
LR_DEV DirectionSample ds_from_si(const DScene &sc, const SurfaceInteraction &si, V3 ref_p) {
    DirectionSample ds; ds.p = si.p; ds.n = si.sh.n; ds.uv = si.uv; ds.pdf = 0.f; ds.delta = false;
    V3 rel = si.p - ref_p; ds.dist = norm(rel);
    ds.d = si_valid(si) ? rel / ds.dist : -si.wi;
    ds.emitter = si_valid(si) ? sc.meshes[si.mesh].emitter : sc.env;
    return ds;
}

} // namespace lr
