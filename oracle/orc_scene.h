/*
 * TEST INFRASTRUCTURE ONLY (see orc_math.h).  Scene container, brute-force-safe
 * BVH, Moeller-Trumbore and surface-interaction reconstruction restated from the
 * reference.
 */
#pragma once
#include "orc_math.h"
#include "../include/lr_api.h"
#include <vector>
#include <memory>
#include <cstdio>

namespace orc {

struct Ray { V3 o, d; float maxt; };
inline V3 ray_at(const Ray &r, float t) { return fma3(r.d, t, r.o); }

struct Tri { V3 p0, p1, p2; uint32_t mesh, face; };

struct Mesh {
    std::vector<V3> pos, nrm; std::vector<V2> uv; std::vector<uint32_t> idx;
    int bsdf, interior, exterior, emitter; bool flip;
    uint32_t tri_offset;       /* first global triangle id */
    float area; std::vector<float> area_cdf;
    bool is_medium_transition() const { return interior >= 0 || exterior >= 0; }
};

/* Interaction3f / SurfaceInteraction3f subset (include/mitsuba/render/interaction.h) */
struct Interaction {
    float t = Inf; V3 p, n;
    bool is_valid() const { return t != Inf; }
    /* interaction.h:165-169 */
    V3 offset_p(V3 d) const {
        float mag = (1.f + std::max(fabsf(p.x), std::max(fabsf(p.y), fabsf(p.z)))) * RayEpsilon;
        mag = mulsign(mag, dot(n, d));
        return fma3(V3(mag), n, p);
    }
    /* interaction.h:140-142 */
    Ray spawn_ray(V3 d) const { return {offset_p(d), d, Largest}; }
    /* interaction.h:145-152 */
    Ray spawn_ray_to(V3 t_) const {
        V3 o = offset_p(t_ - p); V3 d = t_ - o; float dist = norm(d); d = d / dist;
        return {o, d, dist * (1.f - ShadowEpsilon)};
    }
};
struct SurfaceInteraction : Interaction {
    Frame sh; V3 wi, dp_du, dp_dv; V2 uv; int mesh = -1; uint32_t prim = 0;
    V3 to_world(V3 v) const { return sh.to_world(v); }
    V3 to_local(V3 v) const { return sh.to_local(v); }
};
struct Hit { float t = Inf; uint32_t prim = 0xffffffffu; float u = 0, v = 0; bool valid() const { return t != Inf; } };

/* Mesh::moeller_trumbore (include/mitsuba/render/mesh.h:504-526) */
inline bool moeller_trumbore(const Ray &ray, V3 p0, V3 p1, V3 p2, float &t, float &u, float &v) {
    V3 e1 = p1 - p0, e2 = p2 - p0;
    V3 pvec = cross(ray.d, e2);
    float inv_det = 1.f / dot(e1, pvec);
    V3 tvec = ray.o - p0;
    u = dot(tvec, pvec) * inv_det;
    bool active = u >= 0.f && u <= 1.f;
    V3 qvec = cross(tvec, e1);
    v = dot(ray.d, qvec) * inv_det;
    active &= v >= 0.f && u + v <= 1.f;
    t = dot(e2, qvec) * inv_det;
    active &= t >= 0.f && t <= ray.maxt;
    return active;
}

struct BVHNode { V3 lo, hi; int left, right; uint32_t first, count; };

struct Scene;

struct BVH {
    std::vector<BVHNode> nodes; std::vector<uint32_t> order;
    const std::vector<Tri> *tris = nullptr;
    void build(const std::vector<Tri> &t) {
        tris = &t; order.resize(t.size());
        for (size_t i = 0; i < t.size(); ++i) order[i] = (uint32_t) i;
        nodes.clear(); nodes.reserve(2 * t.size() + 1);
        if (!t.empty()) build_rec(0, (uint32_t) t.size());
    }
    int build_rec(uint32_t first, uint32_t count) {
        int id = (int) nodes.size(); nodes.push_back({});
        V3 lo(Inf), hi(-Inf), clo(Inf), chi(-Inf);
        for (uint32_t i = first; i < first + count; ++i) {
            const Tri &tr = (*tris)[order[i]];
            for (const V3 *p : {&tr.p0, &tr.p1, &tr.p2})
                for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], (*p)[k]); hi[k] = std::max(hi[k], (*p)[k]); }
            V3 c = (tr.p0 + tr.p1 + tr.p2) * (1.f / 3.f);
            for (int k = 0; k < 3; ++k) { clo[k] = std::min(clo[k], c[k]); chi[k] = std::max(chi[k], c[k]); }
        }
        nodes[id].lo = lo; nodes[id].hi = hi; nodes[id].first = first; nodes[id].count = count; nodes[id].left = nodes[id].right = -1;
        if (count <= 4) return id;
        V3 ext = chi - clo; int ax = ext.x > ext.y ? (ext.x > ext.z ? 0 : 2) : (ext.y > ext.z ? 1 : 2);
        uint32_t mid = first + count / 2;
        std::nth_element(order.begin() + first, order.begin() + mid, order.begin() + first + count,
            [&](uint32_t a, uint32_t b) {
                const Tri &A = (*tris)[a], &B = (*tris)[b];
                return (A.p0[ax] + A.p1[ax] + A.p2[ax]) < (B.p0[ax] + B.p1[ax] + B.p2[ax]);
            });
        int l = build_rec(first, mid - first); int r = build_rec(mid, first + count - mid);
        nodes[id].left = l; nodes[id].right = r; nodes[id].count = 0;
        return id;
    }
    static bool slab(const BVHNode &n, const Ray &r, V3 inv, float tmax) {
        /* conservative: widened by a relative epsilon so the exact triangle test decides */
        float t0 = 0.f, t1 = tmax;
        for (int k = 0; k < 3; ++k) {
            float a = (n.lo[k] - r.o[k]) * inv[k], b = (n.hi[k] - r.o[k]) * inv[k];
            if (a > b) std::swap(a, b);
            if (std::isnan(a) || std::isnan(b)) continue;
            a -= fabsf(a) * 4e-7f + 1e-30f; b += fabsf(b) * 4e-7f + 1e-30f;
            t0 = std::max(t0, a); t1 = std::min(t1, b);
        }
        return t0 <= t1;
    }
    /* closest hit; ties resolved towards the smaller global primitive index */
    Hit intersect(const Ray &ray) const {
        Hit h; if (nodes.empty()) return h;
        V3 inv(1.f / ray.d.x, 1.f / ray.d.y, 1.f / ray.d.z);
        int stack[128]; int sp = 0; stack[sp++] = 0; float tmax = ray.maxt;
        while (sp) {
            const BVHNode &n = nodes[stack[--sp]];
            if (!slab(n, ray, inv, tmax)) continue;
            if (n.left < 0) {
                for (uint32_t i = n.first; i < n.first + n.count; ++i) {
                    const Tri &tr = (*tris)[order[i]]; float t, u, v;
                    Ray r2 = ray; r2.maxt = tmax;
                    if (moeller_trumbore(r2, tr.p0, tr.p1, tr.p2, t, u, v)) {
                        if (t < h.t || (t == h.t && order[i] < h.prim)) { h.t = t; h.prim = order[i]; h.u = u; h.v = v; tmax = t; }
                    }
                }
            } else { stack[sp++] = n.left; stack[sp++] = n.right; }
        }
        return h;
    }
    bool occluded(const Ray &ray) const { return intersect(ray).valid(); }
};

struct Grid {
    std::vector<float> data; uint32_t res[3]; uint32_t ch; M4 to_local; V3 bmin, bmax;
    uint32_t filter, wrap; float maxv; bool differentiable; bool fixed_max = false;
    std::vector<double> grad;     /* adjoint accumulator (prbvolpath) */
    int wrap_i(int p, int r) const {
        if (wrap == LR_WRAP_CLAMP) return std::min(std::max(p, 0), r - 1);
        /* texture.h:1369-1395 (repeat / mirror) */
        int sh = p < 0 ? p + 1 : p; int div = sh / r;
        int mod = p - div * r; if (mod < 0) mod += r;
        if (wrap == LR_WRAP_MIRROR) { bool keep = ((div & 1) == 0) != (p < 0); if (!keep) mod = r - 1 - mod; }
        return mod;
    }
    /* dr::Texture<Float,3>::eval_nonaccel (ext/drjit/include/drjit/texture.h:484-556)
       via GridVolume::interpolate_1/3 (src/volumes/grid.cpp:553-583).
       Returns the 8 corner indices / weights when asked (adjoint). */
    void eval(V3 pw, float *out, uint32_t *cidx = nullptr, float *cw = nullptr) const {
        V3 p = xform_point_affine(to_local, pw);
        if (filter == LR_FILTER_NEAREST) {
            int ix = wrap_i((int) floorf(p.x * res[0]), res[0]), iy = wrap_i((int) floorf(p.y * res[1]), res[1]),
                iz = wrap_i((int) floorf(p.z * res[2]), res[2]);
            size_t idx = ((size_t) iz * res[1] + iy) * res[0] + ix;
            for (uint32_t c = 0; c < ch; ++c) out[c] = data[idx * ch + c];
            if (cidx) { for (int k = 0; k < 8; ++k) { cidx[k] = (uint32_t) idx; cw[k] = k == 0 ? 1.f : 0.f; } }
            return;
        }
        float fx = fmaf(p.x, (float) res[0], -.5f), fy = fmaf(p.y, (float) res[1], -.5f), fz = fmaf(p.z, (float) res[2], -.5f);
        int ix = (int) floorf(fx), iy = (int) floorf(fy), iz = (int) floorf(fz);
        float w1x = fx - ix, w1y = fy - iy, w1z = fz - iz, w0x = 1.f - w1x, w0y = 1.f - w1y, w0z = 1.f - w1z;
        int x0 = wrap_i(ix, res[0]), x1 = wrap_i(ix + 1, res[0]), y0 = wrap_i(iy, res[1]), y1 = wrap_i(iy + 1, res[1]),
            z0 = wrap_i(iz, res[2]), z1 = wrap_i(iz + 1, res[2]);
        const int xs[2] = {x0, x1}, ys[2] = {y0, y1}, zs[2] = {z0, z1};
        const float wx[2] = {w0x, w1x}, wy[2] = {w0y, w1y}, wz[2] = {w0z, w1z};
        for (uint32_t c = 0; c < ch; ++c) out[c] = 0.f;
        for (int k = 0; k < 8; ++k) {
            int a = k & 1, b = (k >> 1) & 1, cc = (k >> 2) & 1;
            size_t idx = ((size_t) zs[cc] * res[1] + ys[b]) * res[0] + xs[a];
            float w = wx[a] * wy[b] * wz[cc];
            for (uint32_t c = 0; c < ch; ++c) out[c] = fmaf(data[idx * ch + c], w, out[c]);
            if (cidx) { cidx[k] = (uint32_t) idx; cw[k] = w; }
        }
    }
    V3 eval3(V3 pw) const { float o[3]; eval(pw, o); return ch == 1 ? V3(o[0]) : V3(o[0], o[1], o[2]); }
    void recompute_max() { maxv = 0.f; for (float v : data) maxv = std::max(maxv, v); }
};

/* `bitmap` texture (src/textures/bitmap.cpp) on dr::Texture<Float,2>::eval_nonaccel /
   eval_fetch_nonaccel (ext/drjit/include/drjit/texture.h:484-556,704-735; wrap :1369-1396) */
struct Texture2D {
    std::vector<float> data; int w = 0, h = 0, ch = 1; uint32_t filter = 0, wrap = 0; float to_uv[6] = {1, 0, 0, 0, 1, 0};
    static float lum(const float *c) { return c[0] * 0.212671f + c[1] * 0.715160f + c[2] * 0.072169f; }
    int wrap_i(int p, int r) const {
        if (wrap == LR_WRAP_CLAMP) return std::min(std::max(p, 0), r - 1);
        int sh = p < 0 ? p + 1 : p; int div = sh / r;
        int mod = p - div * r; if (mod < 0) mod += r;
        if (wrap == LR_WRAP_MIRROR) { bool keep = ((div & 1) == 0) != (p < 0); if (!keep) mod = r - 1 - mod; }
        return mod;
    }
    /* m_transform * si.uv (AffineTransform3f on a Point2f: fmadd chain of Matrix * point) */
    V2 xform(V2 uv) const {
        return {fmaf(to_uv[0], uv.x, fmaf(to_uv[1], uv.y, to_uv[2])), fmaf(to_uv[3], uv.x, fmaf(to_uv[4], uv.y, to_uv[5]))};
    }
    /* the four bilinear taps (x-fastest order 00,10,01,11) and the fractional weights */
    void fetch4(V2 uv, const float *tap[4], float &w1x, float &w1y) const {
        float fx = fmaf(uv.x, (float) w, -.5f), fy = fmaf(uv.y, (float) h, -.5f);
        int ix = (int) floorf(fx), iy = (int) floorf(fy);
        w1x = fx - (float) ix; w1y = fy - (float) iy;
        int x0 = wrap_i(ix, w), x1 = wrap_i(ix + 1, w), y0 = wrap_i(iy, h), y1 = wrap_i(iy + 1, h);
        tap[0] = &data[((size_t) y0 * w + x0) * ch]; tap[1] = &data[((size_t) y0 * w + x1) * ch];
        tap[2] = &data[((size_t) y1 * w + x0) * ch]; tap[3] = &data[((size_t) y1 * w + x1) * ch];
    }
    /* interpolate_1 / interpolate_3 (bitmap.cpp:781-813); `out` has `ch` entries */
    void eval_raw(V2 uv_in, float *out) const {
        V2 uv = xform(uv_in);
        if (filter == LR_TEX_NEAREST) {
            int ix = wrap_i((int) floorf(uv.x * (float) w), w), iy = wrap_i((int) floorf(uv.y * (float) h), h);
            for (int c = 0; c < ch; ++c) out[c] = data[((size_t) iy * w + ix) * ch + c];
            return;
        }
        const float *t[4]; float w1x, w1y; fetch4(uv, t, w1x, w1y);
        float w0x = 1.f - w1x, w0y = 1.f - w1y;
        const float wt[4] = {w0x * w0y, w1x * w0y, w0x * w1y, w1x * w1y};
        for (int c = 0; c < ch; ++c) out[c] = 0.f;
        for (int k = 0; k < 4; ++k) for (int c = 0; c < ch; ++c) out[c] = fmaf(t[k][c], wt[k], out[c]);
    }
    /* Texture::eval for *_rgb variants (bitmap.cpp:443-485) */
    V3 eval(V2 uv) const { float o[3]; eval_raw(uv, o); return ch == 1 ? V3(o[0]) : V3(o[0], o[1], o[2]); }
    /* eval_1 (:487-507) */
    float eval_1(V2 uv) const { float o[3]; eval_raw(uv, o); return ch == 1 ? o[0] : lum(o); }
    /* eval_1_grad (:509-578) */
    V2 eval_1_grad(V2 uv_in) const {
        if (filter != LR_TEX_BILINEAR) return {0.f, 0.f};
        V2 uv = xform(uv_in);
        const float *t[4]; float w1x, w1y; fetch4(uv, t, w1x, w1y);
        float f00, f10, f01, f11;
        if (ch == 1) { f00 = t[0][0]; f10 = t[1][0]; f01 = t[2][0]; f11 = t[3][0]; }
        else { f00 = lum(t[0]); f10 = lum(t[1]); f01 = lum(t[2]); f11 = lum(t[3]); }
        float w0x = 1.f - w1x, w0y = 1.f - w1y;
        float dx = fmaf(w0y, f10 - f00, w1y * (f11 - f01)), dy = fmaf(w0x, f01 - f00, w1x * (f11 - f10));
        /* transpose multiply by the uv transform, then scale by the resolution */
        float du = to_uv[0] * dx + to_uv[3] * dy, dv = to_uv[1] * dx + to_uv[4] * dy;
        return {(float) w * du, (float) h * dv};
    }
};

struct Medium {
    lr_medium m; float max_density = 0.f;     /* heterogeneous: scale * sigma_t->max() (heterogeneous.cpp:163) */
};

struct Hier2D;   /* envmap warp, orc_emitter.h */

struct Emitter {
    lr_emitter e;
    std::vector<float> pixels;      /* envmap, (W+1) x H x 3 incl. duplicated column */
    std::shared_ptr<Hier2D> warp;
    M4 to_world, to_local;
};

struct Scene {
    std::vector<Mesh> meshes; std::vector<Tri> tris; BVH bvh;
    std::vector<lr_bsdf> bsdfs; std::vector<Texture2D> textures; std::vector<Medium> media; std::vector<Grid> grids; std::vector<Emitter> emitters;
    lr_sensor sensor; lr_integrator integ;
    int env = -1;                   /* index of the environment emitter */
    V3 bs_center; float bs_radius = 0.f;   /* scene bounding sphere (Scene bbox, src/render/scene.cpp) */
    float gauss_coeff[10];          /* GaussianFilter::m_coeff (src/rfilters/gaussian.cpp:60-96) */
    M4 cam_to_world, sample_to_camera;

    Hit ray_intersect_prelim(const Ray &r) const { return bvh.intersect(r); }
    bool ray_test(const Ray &r) const { return bvh.occluded(r); }

    /* Mesh::compute_surface_interaction (src/render/mesh.cpp:1488-1650) +
       finalize_surface_interaction (interaction.h:516-536) + initialize_sh_frame (:290-300) */
    SurfaceInteraction compute_si(const Ray &ray, const Hit &h) const {
        SurfaceInteraction si;
        if (!h.valid()) { si.t = Inf; si.wi = -ray.d; si.mesh = -1; return si; }
        const Tri &tr = tris[h.prim]; const Mesh &me = meshes[tr.mesh];
        uint32_t i0 = me.idx[3 * tr.face], i1 = me.idx[3 * tr.face + 1], i2 = me.idx[3 * tr.face + 2];
        V3 p0 = tr.p0, p1 = tr.p1, p2 = tr.p2;
        float b1 = h.u, b2 = h.v, b0 = 1.f - b1 - b2;
        si.p = fma3(p0, b0, fma3(p1, b1, p2 * b2));
        si.t = h.t; si.mesh = (int) tr.mesh; si.prim = h.prim;
        si.n = normalize(cross(p1 - p0, p2 - p0));
        si.uv = {b1, b2};
        coordinate_system(si.n, si.dp_du, si.dp_dv);
        V3 dp0 = p1 - p0, dp1 = p2 - p0;
        if (!me.uv.empty()) {
            V2 uv0 = me.uv[i0], uv1 = me.uv[i1], uv2 = me.uv[i2];
            si.uv = {fmaf(uv2.x, b2, fmaf(uv1.x, b1, uv0.x * b0)), fmaf(uv2.y, b2, fmaf(uv1.y, b1, uv0.y * b0))};
            V2 duv0 = {uv1.x - uv0.x, uv1.y - uv0.y}, duv1 = {uv2.x - uv0.x, uv2.y - uv0.y};
            float det = fmaf(duv0.x, duv1.y, -(duv0.y * duv1.x)), inv_det = 1.f / det;
            if (det != 0.f) {
                si.dp_du = V3(fmaf(duv1.y, dp0.x, -(duv0.y * dp1.x)), fmaf(duv1.y, dp0.y, -(duv0.y * dp1.y)), fmaf(duv1.y, dp0.z, -(duv0.y * dp1.z))) * inv_det;
                si.dp_dv = V3(fmaf(-duv1.x, dp0.x, duv0.x * dp1.x), fmaf(-duv1.x, dp0.y, duv0.x * dp1.y), fmaf(-duv1.x, dp0.z, duv0.x * dp1.z)) * inv_det;
            }
        }
        if (!me.nrm.empty()) {
            V3 n0 = me.nrm[i0], n1 = me.nrm[i1], n2 = me.nrm[i2];
            V3 n = fma3(n2, b2, fma3(n1, b1, n0 * b0));
            float il = 1.f / sqrtf(sqnorm(n));
            si.sh.n = n * il;
        } else si.sh.n = si.n;
        if (me.flip) { si.n = -si.n; si.sh.n = -si.sh.n; }                 /* mesh.cpp:1652-1655 */
        /* initialize_sh_frame: Gram-Schmidt against dp_du */
        si.sh.s = normalize(fma3(si.sh.n, -dot(si.sh.n, si.dp_du), si.dp_du));
        if (si.dp_du.x == 0.f && si.dp_du.y == 0.f && si.dp_du.z == 0.f) { V3 t_; coordinate_system(si.sh.n, si.sh.s, t_); }
        si.sh.t = cross(si.sh.n, si.sh.s);
        si.wi = si.to_local(-ray.d);
        return si;
    }
    SurfaceInteraction ray_intersect(const Ray &r) const { return compute_si(r, bvh.intersect(r)); }

    /* si.emitter(scene): shape emitter, or the environment when the ray escaped */
    int emitter_of(const SurfaceInteraction &si) const { return si.is_valid() ? meshes[si.mesh].emitter : env; }
    /* interaction.h:322-345 */
    int target_medium(const SurfaceInteraction &si, V3 d) const {
        const Mesh &m = meshes[si.mesh];
        return dot(d, si.n) > 0.f ? m.exterior : m.interior;
    }
};

} // namespace orc
