/*
 * TEST INFRASTRUCTURE ONLY (see orc_math.h).  Emitters (area / constant /
 * envmap with the hierarchical 2D warp) restated from the reference.
 */
#pragma once
#include "orc_scene.h"
#include "orc_bsdf.h"

namespace orc {

struct DirectionSample {
    V3 p, n, d; V2 uv; float pdf = 0.f, dist = 0.f; bool delta = false; int emitter = -1;
};

/* Hierarchical2D<Float, 0> (include/mitsuba/core/distr_2d.h:371-830) */
struct Hier2D {
    struct Level {
        uint32_t size, width; std::vector<float> data;
        Level(uint32_t w, uint32_t h) : size(w * h), width(w), data((size_t) w * h, 0.f) {}
        /* 2x2 patches contiguous (distr_2d.h Level::index) */
        uint32_t index(uint32_t x, uint32_t y) const { return ((x & 1u) | (((x & ~1u) | (y & 1u)) << 1)) + ((y & ~1u) * width); }
    };
    std::vector<Level> levels; V2 patch_size, inv_patch_size; uint32_t max_patch_x, max_patch_y;

    static uint32_t log2i_ceil(uint32_t v) { uint32_t r = 0; while ((1u << r) < v) ++r; return r; }

    Hier2D(const float *data, uint32_t w, uint32_t h) {
        uint32_t npx = w - 1, npy = h - 1;
        patch_size = {1.f / npx, 1.f / npy}; inv_patch_size = {(float) npx, (float) npy};
        uint32_t max_level = log2i_ceil(std::max(npx, npy));
        max_patch_x = npx - 1; max_patch_y = npy - 1;
        levels.reserve(max_level + 2);
        levels.emplace_back(w, h);
        uint32_t lx = npx, ly = npy;
        for (int level = (int) max_level; level >= 0; --level) {
            lx += lx & 1u; ly += ly & 1u;
            levels.emplace_back(lx, ly);
            lx >>= 1; ly >>= 1;
        }
        const float *in = data; double sum = 0.0;
        for (uint32_t y = 0; y < npy; ++y) {
            for (uint32_t x = 0; x < npx; ++x) {
                float avg = .25f * (in[0] + in[1] + in[w] + in[w + 1]);
                sum += (double) avg;
                levels[1].data[levels[1].index(x, y)] = avg;
                ++in;
            }
            ++in;
        }
        float scale = (float) ((double) (npx * npy) / sum);
        for (uint32_t i = 0; i < levels[0].size; ++i) levels[0].data[i] = data[i] * scale;
        for (uint32_t i = 0; i < levels[1].size; ++i) levels[1].data[i] *= scale;
        lx = npx; ly = npy;
        for (uint32_t level = 2; level <= max_level + 1; ++level) {
            const Level &l0 = levels[level - 1]; Level &l1 = levels[level];
            lx = (lx + 1u) >> 1; ly = (ly + 1u) >> 1;
            for (uint32_t y = 0; y < ly; ++y)
                for (uint32_t x = 0; x < lx; ++x) {
                    const float *d0 = l0.data.data() + l0.index(x * 2, y * 2);
                    l1.data[l1.index(x, y)] = d0[0] + d0[1] + d0[2] + d0[3];
                }
        }
    }
    static float clip01(float v) { return std::max(std::min(v, 1.f), 0.f); }
    /* warp.h:446-453 */
    static float interval_to_linear(float v0, float v1, float s) {
        return fabsf(v0 - v1) > 1e-4f * (v0 + v1) ? (v0 - safe_sqrt(lerp(sqr(v0), sqr(v1), s))) / (v0 - v1) : s;
    }
    /* distr_2d.h sample() */
    V2 sample(V2 s, float &pdf) const {
        s.x = clip01(s.x); s.y = clip01(s.y);
        uint32_t ox = 0, oy = 0;
        for (int l = (int) levels.size() - 2; l > 0; --l) {
            const Level &lv = levels[l];
            ox <<= 1; oy <<= 1;
            uint32_t oi = lv.index(ox, oy);
            float v00 = lv.data[oi], v10 = lv.data[oi + 1], v01 = lv.data[oi + 2], v11 = lv.data[oi + 3];
            s.x = clip01(s.x); s.y = clip01(s.y);
            float r0 = v00 + v10, r1 = v01 + v11;
            s.y *= r0 + r1;
            bool mask = s.y > r0;
            if (mask) { oy += 1u; s.y -= r0; }
            s.y /= mask ? r1 : r0;
            float c0 = mask ? v01 : v00, c1 = mask ? v11 : v10;
            s.x *= c0 + c1;
            mask = s.x > c0;
            if (mask) s.x -= c0;
            s.x /= mask ? c1 : c0;
            if (mask) ox += 1u;
        }
        const Level &l0 = levels[0];
        uint32_t oi = ox + oy * l0.width;
        float v00 = l0.data[oi], v10 = l0.data[oi + 1], v01 = l0.data[oi + l0.width], v11 = l0.data[oi + l0.width + 1];
        /* warp::square_to_bilinear (warp.h:480-494) */
        float r0 = v00 + v10, r1 = v01 + v11;
        s.y = interval_to_linear(r0, r1, s.y);
        float c0 = lerp(v00, v01, s.y), c1 = lerp(v10, v11, s.y);
        s.x = interval_to_linear(c0, c1, s.x);
        pdf = lerp(c0, c1, s.x);
        return {((float) (int) ox + s.x) * patch_size.x, ((float) (int) oy + s.y) * patch_size.y};
    }
    float eval(V2 pos) const {
        pos.x = clip01(pos.x) * inv_patch_size.x; pos.y = clip01(pos.y) * inv_patch_size.y;
        uint32_t ox = std::min((uint32_t) (int) pos.x, max_patch_x), oy = std::min((uint32_t) (int) pos.y, max_patch_y);
        pos.x -= (float) (int) ox; pos.y -= (float) (int) oy;
        const Level &l0 = levels[0]; uint32_t oi = ox + oy * l0.width;
        float v00 = l0.data[oi], v10 = l0.data[oi + 1], v01 = l0.data[oi + l0.width], v11 = l0.data[oi + l0.width + 1];
        return lerp(lerp(v00, v10, pos.x), lerp(v01, v11, pos.x), pos.y);
    }
};

inline M4 invert_affine(const M4 &A) {
    /* general 4x4 inverse in double (host-side set-up, not on the per-sample path) */
    double m[16], inv[16]; for (int i = 0; i < 16; ++i) m[i] = A.m[i];
    inv[0] = m[5]*m[10]*m[15]-m[5]*m[11]*m[14]-m[9]*m[6]*m[15]+m[9]*m[7]*m[14]+m[13]*m[6]*m[11]-m[13]*m[7]*m[10];
    inv[4] = -m[4]*m[10]*m[15]+m[4]*m[11]*m[14]+m[8]*m[6]*m[15]-m[8]*m[7]*m[14]-m[12]*m[6]*m[11]+m[12]*m[7]*m[10];
    inv[8] = m[4]*m[9]*m[15]-m[4]*m[11]*m[13]-m[8]*m[5]*m[15]+m[8]*m[7]*m[13]+m[12]*m[5]*m[11]-m[12]*m[7]*m[9];
    inv[12] = -m[4]*m[9]*m[14]+m[4]*m[10]*m[13]+m[8]*m[5]*m[14]-m[8]*m[6]*m[13]-m[12]*m[5]*m[10]+m[12]*m[6]*m[9];
    inv[1] = -m[1]*m[10]*m[15]+m[1]*m[11]*m[14]+m[9]*m[2]*m[15]-m[9]*m[3]*m[14]-m[13]*m[2]*m[11]+m[13]*m[3]*m[10];
    inv[5] = m[0]*m[10]*m[15]-m[0]*m[11]*m[14]-m[8]*m[2]*m[15]+m[8]*m[3]*m[14]+m[12]*m[2]*m[11]-m[12]*m[3]*m[10];
    inv[9] = -m[0]*m[9]*m[15]+m[0]*m[11]*m[13]+m[8]*m[1]*m[15]-m[8]*m[3]*m[13]-m[12]*m[1]*m[11]+m[12]*m[3]*m[9];
    inv[13] = m[0]*m[9]*m[14]-m[0]*m[10]*m[13]-m[8]*m[1]*m[14]+m[8]*m[2]*m[13]+m[12]*m[1]*m[10]-m[12]*m[2]*m[9];
    inv[2] = m[1]*m[6]*m[15]-m[1]*m[7]*m[14]-m[5]*m[2]*m[15]+m[5]*m[3]*m[14]+m[13]*m[2]*m[7]-m[13]*m[3]*m[6];
    inv[6] = -m[0]*m[6]*m[15]+m[0]*m[7]*m[14]+m[4]*m[2]*m[15]-m[4]*m[3]*m[14]-m[12]*m[2]*m[7]+m[12]*m[3]*m[6];
    inv[10] = m[0]*m[5]*m[15]-m[0]*m[7]*m[13]-m[4]*m[1]*m[15]+m[4]*m[3]*m[13]+m[12]*m[1]*m[7]-m[12]*m[3]*m[5];
    inv[14] = -m[0]*m[5]*m[14]+m[0]*m[6]*m[13]+m[4]*m[1]*m[14]-m[4]*m[2]*m[13]-m[12]*m[1]*m[6]+m[12]*m[2]*m[5];
    inv[3] = -m[1]*m[6]*m[11]+m[1]*m[7]*m[10]+m[5]*m[2]*m[11]-m[5]*m[3]*m[10]-m[9]*m[2]*m[7]+m[9]*m[3]*m[6];
    inv[7] = m[0]*m[6]*m[11]-m[0]*m[7]*m[10]-m[4]*m[2]*m[11]+m[4]*m[3]*m[10]+m[8]*m[2]*m[7]-m[8]*m[3]*m[6];
    inv[11] = -m[0]*m[5]*m[11]+m[0]*m[7]*m[9]+m[4]*m[1]*m[11]-m[4]*m[3]*m[9]-m[8]*m[1]*m[7]+m[8]*m[3]*m[5];
    inv[15] = m[0]*m[5]*m[10]-m[0]*m[6]*m[9]-m[4]*m[1]*m[10]+m[4]*m[2]*m[9]+m[8]*m[1]*m[6]-m[8]*m[2]*m[5];
    double det = m[0]*inv[0]+m[1]*inv[4]+m[2]*inv[8]+m[3]*inv[12];
    M4 R; for (int i = 0; i < 16; ++i) R.m[i] = (float) (inv[i] / det);
    return R;
}

/* envmap set-up (src/emitters/envmap.cpp:141-232): append the periodic column,
   importance map lum*sin(theta), hierarchical warp */
inline void envmap_prepare(Emitter &em) {
    const lr_emitter &e = em.e; uint32_t W = e.width, H = e.height, W1 = W + 1;
    em.pixels.assign((size_t) W1 * H * 3, 0.f);
    std::vector<float> lum((size_t) W1 * H);
    float theta_scale = 1.f / (H - 1) * Pi;
    for (uint32_t y = 0; y < H; ++y) {
        float sin_theta = sinf(y * theta_scale);
        for (uint32_t x = 0; x < W; ++x) {
            const float *px = e.pixels + ((size_t) y * W + x) * 3;
            float l = luminance(V3(px[0], px[1], px[2])); l = std::max(l, 0.f);
            lum[(size_t) y * W1 + x] = l * sin_theta;
            for (int c = 0; c < 3; ++c) em.pixels[((size_t) y * W1 + x) * 3 + c] = px[c];
        }
        lum[(size_t) y * W1 + W] = lum[(size_t) y * W1];
        for (int c = 0; c < 3; ++c) em.pixels[((size_t) y * W1 + W) * 3 + c] = em.pixels[((size_t) y * W1) * 3 + c];
    }
    em.warp = std::make_shared<Hier2D>(lum.data(), W1, H);
    memcpy(em.to_world.m, e.to_world, sizeof(float) * 16);
    em.to_local = invert_affine(em.to_world);
}

struct EmitterOps {
    const Scene &sc;
    explicit EmitterOps(const Scene &s) : sc(s) {}

    float bs_radius_env() const { return std::max(RayEpsilon, sc.bs_radius * (1.f + RayEpsilon)); }

    /* envmap.cpp:527-580 eval_spectrum (rgb branch) */
    V3 env_eval_uv(const Emitter &em, V2 uv) const {
        uint32_t rx = em.e.width + 1, ry = em.e.height;
        uv.x -= .5f / (rx - 1u);
        uv.x -= floorf(uv.x); uv.y -= floorf(uv.y);
        uv.x *= (float) (rx - 1u); uv.y *= (float) (ry - 1u);
        uint32_t px = std::min((uint32_t) uv.x, rx - 2u), py = std::min((uint32_t) uv.y, ry - 2u);
        float w1x = uv.x - (float) px, w1y = uv.y - (float) py, w0x = 1.f - w1x, w0y = 1.f - w1y;
        size_t index = (size_t) py * rx + px; const float *d = em.pixels.data();
        V3 r;
        for (int c = 0; c < 3; ++c) {
            float v00 = d[index * 3 + c], v10 = d[(index + 1) * 3 + c], v01 = d[(index + rx) * 3 + c], v11 = d[(index + rx + 1) * 3 + c];
            float v0 = fmaf(w0x, v00, w1x * v10), v1 = fmaf(w0x, v01, w1x * v11);
            r[c] = fmaf(w0y, v0, w1y * v1) * em.e.scale;
        }
        return r;
    }
    static float safe_acos(float x) { return acosf(std::min(std::max(x, -1.f), 1.f)); }

    /* Emitter::eval(si): `wi_world_neg` = -si.wi in world space = direction of travel for escaped rays;
       for area lights cos_theta(si.wi) > 0 (area.cpp:83-90) */
    V3 eval(int id, const SurfaceInteraction &si) const {
        const Emitter &em = sc.emitters[id]; const lr_emitter &e = em.e;
        switch (e.kind) {
            case LR_EMITTER_AREA: return si.wi.z > 0.f ? V3(e.radiance[0], e.radiance[1], e.radiance[2]) : V3(0.f);
            case LR_EMITTER_CONSTANT: return V3(e.radiance[0], e.radiance[1], e.radiance[2]);
            case LR_EMITTER_ENVMAP: {                /* envmap.cpp:353-363; escaped rays: si.wi = -ray.d in world space */
                V3 v = xform_vector(em.to_local, -si.wi);
                V2 uv = {atan2f(v.x, -v.z) * InvTwoPi, safe_acos(v.y) * InvPi};
                return env_eval_uv(em, uv);
            }
        }
        return V3(0.f);
    }

    /* Emitter::sample_direction */
    DirectionSample sample_direction(int id, const Interaction &it, V2 u, V3 &spec) const {
        const Emitter &em = sc.emitters[id]; const lr_emitter &e = em.e; DirectionSample ds; ds.emitter = id; spec = V3(0.f);
        switch (e.kind) {
        case LR_EMITTER_AREA: {                      /* area.cpp:118-168, shape.cpp:344-361, rectangle.cpp:181-199 */
            bool active = true;
            if (e.is_rectangle) {
                float a = fmaf(u.x, 2.f, -1.f), b = fmaf(u.y, 2.f, -1.f);
                V3 o(e.rect_origin[0], e.rect_origin[1], e.rect_origin[2]), eu(e.rect_edge_u[0], e.rect_edge_u[1], e.rect_edge_u[2]),
                   ev(e.rect_edge_v[0], e.rect_edge_v[1], e.rect_edge_v[2]);
                ds.p = fma3(eu, a, fma3(ev, b, o));
                ds.n = V3(e.rect_normal[0], e.rect_normal[1], e.rect_normal[2]);
                ds.pdf = e.inv_area; ds.uv = u;
            } else {
                /* Mesh::sample_position (src/render/mesh.cpp): area-weighted face pick with sample re-use */
                const Mesh &me = sc.meshes[e.mesh];
                float sy = u.y; uint32_t n = (uint32_t) me.area_cdf.size();
                float target = sy * me.area;
                uint32_t f = (uint32_t) (std::upper_bound(me.area_cdf.begin(), me.area_cdf.end(), target) - me.area_cdf.begin());
                f = std::min(f, n - 1);
                float c0 = f ? me.area_cdf[f - 1] : 0.f, c1 = me.area_cdf[f];
                sy = (target - c0) / (c1 - c0);
                V2 b = square_to_uniform_triangle({u.x, sy});
                V3 p0 = me.pos[me.idx[3 * f]], p1 = me.pos[me.idx[3 * f + 1]], p2 = me.pos[me.idx[3 * f + 2]];
                V3 e0 = p1 - p0, e1 = p2 - p0;
                ds.p = fma3(e0, b.x, fma3(e1, b.y, p0));
                ds.n = normalize(cross(e0, e1)); ds.pdf = 1.f / me.area; ds.uv = b;
            }
            ds.d = ds.p - it.p; float dist2 = sqnorm(ds.d); ds.dist = sqrtf(dist2); ds.d = ds.d / ds.dist;
            float dp = abs_dot(ds.d, ds.n), x = dist2 / dp;
            ds.pdf *= std::isfinite(x) ? x : 0.f;
            active &= dot(ds.d, ds.n) < 0.f && ds.pdf != 0.f;
            if (active) spec = V3(e.radiance[0], e.radiance[1], e.radiance[2]) / ds.pdf;
            return ds;
        }
        case LR_EMITTER_CONSTANT: {                  /* constant.cpp:121-150 */
            V3 d = square_to_uniform_sphere(u);
            float radius = std::max(bs_radius_env(), norm(it.p - sc.bs_center)), dist = 2.f * radius;
            ds.p = fma3(d, dist, it.p); ds.n = -d; ds.uv = u; ds.pdf = InvFourPi; ds.d = d; ds.dist = dist;
            spec = V3(e.radiance[0], e.radiance[1], e.radiance[2]) / ds.pdf;
            return ds;
        }
        case LR_EMITTER_ENVMAP: {                    /* envmap.cpp:415-457 */
            float pdf; V2 uv = em.warp->sample(u, pdf);
            uv.x += .5f / (float) (em.e.width + 1 - 1);
            bool active = pdf > 0.f;
            float theta = uv.y * Pi, phi = uv.x * (2.f * Pi);
            float st, ct, sp, cp; sincos_(theta, st, ct); sincos_(phi, sp, cp);
            V3 d0(cp * st, sp * st, ct); V3 d(d0.y, d0.z, -d0.x);
            float radius = std::max(bs_radius_env(), norm(it.p - sc.bs_center)), dist = 2.f * radius;
            float inv_sin_theta = 1.f / sqrtf(std::max(sqr(d.x) + sqr(d.z), sqr(Epsilon)));
            d = xform_vector(em.to_world, d);
            ds.p = it.p + d * dist; ds.n = -d; ds.uv = uv;
            ds.pdf = active ? pdf * inv_sin_theta * (1.f / (2.f * sqr(Pi))) : 0.f;
            ds.d = d; ds.dist = dist;
            if (active) spec = env_eval_uv(em, uv) / ds.pdf;
            return ds;
        }
        }
        return ds;
    }

    /* Emitter::pdf_direction */
    float pdf_direction(int id, const Interaction &, const DirectionSample &ds) const {
        const Emitter &em = sc.emitters[id]; const lr_emitter &e = em.e;
        switch (e.kind) {
        case LR_EMITTER_AREA: {                      /* area.cpp:170-197, shape.cpp:363-374 */
            float dp = dot(ds.d, ds.n); if (!(dp < 0.f)) return 0.f;
            float pdf = e.is_rectangle ? e.inv_area : 1.f / sc.meshes[e.mesh].area;
            float adp = fabsf(dp);
            pdf *= adp != 0.f ? (ds.dist * ds.dist) / adp : 0.f;
            return pdf;
        }
        case LR_EMITTER_CONSTANT: return InvFourPi;
        case LR_EMITTER_ENVMAP: {                    /* envmap.cpp:459-476 */
            V3 d = xform_vector(em.to_local, ds.d);
            V2 uv = {atan2f(d.x, -d.z) * InvTwoPi, safe_acos(d.y) * InvPi};
            uv.x -= .5f / (float) (em.e.width + 1 - 1u);
            uv.x -= floorf(uv.x); uv.y -= floorf(uv.y);
            float inv_sin_theta = 1.f / sqrtf(std::max(sqr(d.x) + sqr(d.z), sqr(Epsilon)));
            return em.warp->eval(uv) * inv_sin_theta * (1.f / (2.f * sqr(Pi)));
        }
        }
        return 0.f;
    }

    /* Scene::sample_emitter_direction (src/render/scene.cpp:333-384), JIT branch,
       uniform emitter selection (scene.cpp:265-288) */
    DirectionSample sample_emitter_direction(const Interaction &ref, V2 sample, bool test_visibility, V3 &spec) const {
        size_t n = sc.emitters.size(); DirectionSample ds; spec = V3(0.f);
        if (n == 0) return ds;
        uint32_t index = 0; float weight = 1.f;
        if (n > 1) {
            float scaled = sample.x * (float) n;
            index = std::min((uint32_t) scaled, (uint32_t) n - 1u);
            weight = (float) n; sample.x = scaled - (float) index;
        }
        ds = sample_direction((int) index, ref, sample, spec);
        ds.pdf *= 1.f / (float) n;                   /* pdf_emitter: m_emitter_pmf */
        spec *= weight;
        bool active = ds.pdf != 0.f;
        if (test_visibility && active) {
            if (sc.ray_test(ref.spawn_ray_to(ds.p))) { spec = V3(0.f); ds.pdf = 0.f; }
        }
        return ds;
    }
    /* scene.cpp:395-406 */
    float pdf_emitter_direction(const Interaction &ref, const DirectionSample &ds) const {
        if (ds.emitter < 0) return 0.f;
        return pdf_direction(ds.emitter, ref, ds) * (1.f / (float) sc.emitters.size());
    }
    /* DirectionSample(scene, si, ref) (include/mitsuba/render/records.h:173-180) */
    DirectionSample ds_from_si(const SurfaceInteraction &si, const Interaction &ref) const {
        DirectionSample ds; ds.p = si.p; ds.n = si.sh.n; ds.uv = si.uv; ds.pdf = 0.f; ds.delta = false;
        V3 rel = si.p - ref.p; ds.dist = norm(rel);
        /* invalid si: wi holds -ray.d in world space */
        ds.d = si.is_valid() ? rel / ds.dist : -si.wi;
        ds.emitter = sc.emitter_of(si);
        return ds;
    }
};

} // namespace orc
