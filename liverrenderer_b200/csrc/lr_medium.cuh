// lr_medium.cuh -- participating media on the hot path (device code).
// Medium::sample_interaction / transmittance_eval_pdf (src/render/medium.cpp:42-104),
// heterogeneous (src/media/heterogeneous.cpp:155-201), homogeneous (src/media/homogeneous.cpp:137-181),
// gridvolume trilinear lookup (src/volumes/grid.cpp:553-583 ->
// ext/drjit/include/drjit/texture.h:484-556, wrap :1369-1395).
#pragma once
#include "lr_scene.cuh"

namespace lr {

LR_DEV int wrap_i(int p, int r, uint32_t wrap) {
    if (wrap == LR_WRAP_CLAMP) return min(max(p, 0), r - 1);
    int sh = p < 0 ? p + 1 : p; int div = sh / r;
    int mod = p - div * r; if (mod < 0) mod += r;
    if (wrap == LR_WRAP_MIRROR) { bool keep = ((div & 1) == 0) != (p < 0); if (!keep) mod = r - 1 - mod; }
    return mod;
}

// Trilinear footprint of one lookup: base corner indices + weights (kept for the adjoint scatter).
struct GridTap { uint32_t idx[8]; float w[8]; };

// 1- or 3-channel trilinear fetch, corner order 000,100,010,110,001,101,011,111, accumulation
// out = fmadd(v_k, w_k, out) exactly as eval_nonaccel.  Algorithmic traffic: 8 corners x ch x 4 B.
template <bool WANT_TAP>
LR_DEV V3 grid_eval(const DGrid &g, V3 pw, GridTap *tap) {
    V3 p = xform_point(g.to_local, pw);
    if (g.filter == LR_FILTER_NEAREST) {
        int ix = wrap_i((int) floorf(p.x * (float) g.rx), g.rx, g.wrap), iy = wrap_i((int) floorf(p.y * (float) g.ry), g.ry, g.wrap),
            iz = wrap_i((int) floorf(p.z * (float) g.rz), g.rz, g.wrap);
        size_t idx = ((size_t) iz * g.ry + iy) * g.rx + ix;
        if (WANT_TAP) { for (int k = 0; k < 8; ++k) { tap->idx[k] = (uint32_t) idx; tap->w[k] = k == 0 ? 1.f : 0.f; } }
        if (g.ch == 1) return mk3(__ldg(g.data + idx));
        return mk3(__ldg(g.data + idx * 3), __ldg(g.data + idx * 3 + 1), __ldg(g.data + idx * 3 + 2));
    }
    float fx = fmaf(p.x, (float) g.rx, -.5f), fy = fmaf(p.y, (float) g.ry, -.5f), fz = fmaf(p.z, (float) g.rz, -.5f);
    float flx = floorf(fx), fly = floorf(fy), flz = floorf(fz);
    int ix = (int) flx, iy = (int) fly, iz = (int) flz;
    float w1x = fx - flx, w1y = fy - fly, w1z = fz - flz, w0x = 1.f - w1x, w0y = 1.f - w1y, w0z = 1.f - w1z;
    if (!WANT_TAP && g.fat) {                                      // fat-voxel record: same eight values, same weights, same fmadd order
        const int rix = min(max(ix + 1, 0), g.rx), riy = min(max(iy + 1, 0), g.ry), riz = min(max(iz + 1, 0), g.rz);
        const float4 *rec = g.fat + 2 * (((size_t) riz * (g.ry + 1) + riy) * (g.rx + 1) + rix);
        const float4 a = __ldg(rec), b = __ldg(rec + 1);
        float o = 0.f;
        o = fmaf(a.x, __fmul_rn(__fmul_rn(w0x, w0y), w0z), o); o = fmaf(a.y, __fmul_rn(__fmul_rn(w1x, w0y), w0z), o);
        o = fmaf(a.z, __fmul_rn(__fmul_rn(w0x, w1y), w0z), o); o = fmaf(a.w, __fmul_rn(__fmul_rn(w1x, w1y), w0z), o);
        o = fmaf(b.x, __fmul_rn(__fmul_rn(w0x, w0y), w1z), o); o = fmaf(b.y, __fmul_rn(__fmul_rn(w1x, w0y), w1z), o);
        o = fmaf(b.z, __fmul_rn(__fmul_rn(w0x, w1y), w1z), o); o = fmaf(b.w, __fmul_rn(__fmul_rn(w1x, w1y), w1z), o);
        return mk3(o);
    }
    int x0 = wrap_i(ix, g.rx, g.wrap), x1 = wrap_i(ix + 1, g.rx, g.wrap), y0 = wrap_i(iy, g.ry, g.wrap), y1 = wrap_i(iy + 1, g.ry, g.wrap),
        z0 = wrap_i(iz, g.rz, g.wrap), z1 = wrap_i(iz + 1, g.rz, g.wrap);
    uint32_t r00 = ((uint32_t) z0 * g.ry + y0) * g.rx, r10 = ((uint32_t) z0 * g.ry + y1) * g.rx,
             r01 = ((uint32_t) z1 * g.ry + y0) * g.rx, r11 = ((uint32_t) z1 * g.ry + y1) * g.rx;
    uint32_t i[8] = {r00 + x0, r00 + x1, r10 + x0, r10 + x1, r01 + x0, r01 + x1, r11 + x0, r11 + x1};
    float w[8] = {__fmul_rn(__fmul_rn(w0x, w0y), w0z), __fmul_rn(__fmul_rn(w1x, w0y), w0z), __fmul_rn(__fmul_rn(w0x, w1y), w0z), __fmul_rn(__fmul_rn(w1x, w1y), w0z),
                  __fmul_rn(__fmul_rn(w0x, w0y), w1z), __fmul_rn(__fmul_rn(w1x, w0y), w1z), __fmul_rn(__fmul_rn(w0x, w1y), w1z), __fmul_rn(__fmul_rn(w1x, w1y), w1z)};
    if (WANT_TAP) {
#pragma unroll
        for (int k = 0; k < 8; ++k) { tap->idx[k] = i[k]; tap->w[k] = w[k]; }
    }
    if (g.ch == 1) {
        float v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = __ldg(g.data + i[k]);      // 8 independent loads in flight
        float o = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) o = fmaf(v[k], w[k], o);
        return mk3(o);
    }
    V3 o = mk3(0.f);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const float *d = g.data + (size_t) i[k] * 3;
        o.x = fmaf(__ldg(d), w[k], o.x); o.y = fmaf(__ldg(d + 1), w[k], o.y); o.z = fmaf(__ldg(d + 2), w[k], o.z);
    }
    return o;
}

// BoundingBox3f::ray_intersect (include/mitsuba/core/bbox.h:303-340)
LR_DEV bool aabb_intersect(V3 bmin, V3 bmax, const Ray &ray, float &mint, float &maxt) {
    bool active = ray.d.x != 0.f || ray.d.y != 0.f || ray.d.z != 0.f;
    float rx = 1.f / ray.d.x, ry = 1.f / ray.d.y, rz = 1.f / ray.d.z;
    float tminx = ((rx >= 0 ? bmin.x : bmax.x) - ray.o.x) * rx, tmaxx = ((rx >= 0 ? bmax.x : bmin.x) - ray.o.x) * rx;
    float tminy = ((ry >= 0 ? bmin.y : bmax.y) - ray.o.y) * ry, tmaxy = ((ry >= 0 ? bmax.y : bmin.y) - ray.o.y) * ry;
    float tminz = ((rz >= 0 ? bmin.z : bmax.z) - ray.o.z) * rz, tmaxz = ((rz >= 0 ? bmax.z : bmin.z) - ray.o.z) * rz;
    active = active && !((tminx > tmaxy) || (tminy > tmaxx));
    tminx = (tminx > tminy || !isfinite(tminy)) ? tminx : tminy;
    tmaxx = (tmaxx < tmaxy || !isfinite(tmaxy)) ? tmaxx : tmaxy;
    active = active && !((tminx > tmaxz) || (tminz > tmaxx));
    tminx = (tminx > tminz || !isfinite(tminz)) ? tminx : tminz;
    tmaxx = (tmaxx < tmaxz || !isfinite(tmaxz)) ? tmaxx : tmaxz;
    mint = tminx; maxt = tmaxx; return active;
}


// ---- majorant super-grid DDA (production tracking mode, LR_TRACKING_DDA) -----------------------------------
// The sigma_t grid is covered by super-voxels of `cell`^3 voxels; maj[c] bounds the raw grid over the trilinear
// support of super-voxel c (k_majorant_grid).  A ray is walked through the super-voxels in grid-local coordinates
// (same ray parameter t as in world space); inside a super-voxel the free-flight majorant is scale * maj[c].
struct DDA { int ix, iy, iz, sx, sy, sz; float tnx, tny, tnz, tdx, tdy, tdz; };
LR_DEV void dda_init(const DGrid &g, const Ray &ray, float t0, DDA &d) {
    V3 ol = xform_point(g.to_local, ray_at(ray, t0)), dl = xform_vector(g.to_local, ray.d);
    float Sx = (float) g.rx / (float) g.cell, Sy = (float) g.ry / (float) g.cell, Sz = (float) g.rz / (float) g.cell;
    float gx = ol.x * Sx, gy = ol.y * Sy, gz = ol.z * Sz, vx = dl.x * Sx, vy = dl.y * Sy, vz = dl.z * Sz;
    d.ix = min(max((int) floorf(gx), 0), g.mx - 1); d.iy = min(max((int) floorf(gy), 0), g.my - 1); d.iz = min(max((int) floorf(gz), 0), g.mz - 1);
    d.sx = vx > 0.f ? 1 : -1; d.sy = vy > 0.f ? 1 : -1; d.sz = vz > 0.f ? 1 : -1;
    d.tdx = vx != 0.f ? fabsf(1.f / vx) : LR_INF; d.tdy = vy != 0.f ? fabsf(1.f / vy) : LR_INF; d.tdz = vz != 0.f ? fabsf(1.f / vz) : LR_INF;
    d.tnx = vx != 0.f ? t0 + fmaxf(((float) (d.ix + (vx > 0.f ? 1 : 0)) - gx) / vx, 0.f) : LR_INF;
    d.tny = vy != 0.f ? t0 + fmaxf(((float) (d.iy + (vy > 0.f ? 1 : 0)) - gy) / vy, 0.f) : LR_INF;
    d.tnz = vz != 0.f ? t0 + fmaxf(((float) (d.iz + (vz > 0.f ? 1 : 0)) - gz) / vz, 0.f) : LR_INF;
}
LR_DEV float dda_majorant(const DGrid &g, const DDA &d) { return __ldg(g.maj + ((size_t) d.iz * g.my + d.iy) * g.mx + d.ix); }
LR_DEV void dda_advance(const DGrid &g, DDA &d) {
    if (d.tnx <= d.tny && d.tnx <= d.tnz) { d.ix = min(max(d.ix + d.sx, 0), g.mx - 1); d.tnx += d.tdx; }
    else if (d.tny <= d.tnz) { d.iy = min(max(d.iy + d.sy, 0), g.my - 1); d.tny += d.tdy; }
    else { d.iz = min(max(d.iz + d.sz, 0), g.mz - 1); d.tnz += d.tdz; }
}
// Advance the DDA until the carried optical depth `tau` is used up -- tentative collision at `t` under majorant `mu`, returns
// true -- or the end of the segment is reached (false).  Kept as its own loop so that a warp first brings every lane to its
// next collision (cheap steps) and then evaluates the collisions together (expensive: 8 grid taps, RNG, log).
LR_DEV bool dda_walk(const DGrid &g, float mu_scale, DDA &dd, float &t, float &tau, float t_end, float &mu) {
    for (;;) {
        mu = mu_scale * dda_majorant(g, dd);
        const float t_exit = fminf(fminf(dd.tnx, dd.tny), fminf(dd.tnz, t_end));
        const float dtau = mu * (t_exit - t);
        if (dtau >= tau && mu > 0.f) { t = fminf(t + tau / mu, t_exit); return true; }
        tau -= dtau; t = t_exit;
        if (t_exit >= t_end) return false;
        dda_advance(g, dd);
    }
}
// clip of the segment against the volume's AABB exactly as Medium::sample_interaction does (medium.cpp:53-60)
LR_DEV bool medium_clip(const DGrid &g, const Ray &ray, float &mint, float &maxt) {
    bool its = aabb_intersect(g.bmin, g.bmax, ray, mint, maxt);
    its &= (isfinite(mint) || isfinite(maxt));
    if (!its) return false;
    mint = fmaxf(0.f, mint); maxt = fminf(ray.maxt, maxt);
    return true;
}
// The reference weights a free-flight step that ends on a surface (or at the emitter) with Tr / pdf = exp(-majorant * t) / exp(-majorant * t),
// t = the distance from the step's start to that surface, measured with the GLOBAL majorant and NOT clipped to the volume's box
// (volpath.cpp:228-232, :504-511).  The ratio is 1 -- unless exp() underflows to 0 in fp32 (majorant * t > ~103), in which case the
// reference sets the weight to 0.  That happens for a ray that carries a medium through open space (an unbounded segment: si.t = inf,
// or a far emitter), e.g. behind the intersecting shells of the Liver-MultiMesh scene.  The DDA walks reproduce the rule on their last
// step, so that production tracking agrees with the reference on such scenes too.  `t_from`: start of the last step, `t_to`: its end.
LR_DEV bool reference_step_underflows(const DMedium &m, float t_from, float t_to) { return expf(-m.max_density * (t_to - t_from)) == 0.f; }

LR_DEV bool medium_uses_dda(const DScene &sc, const DMedium &m) {
    return sc.integ.tracking == LR_TRACKING_DDA && m.m.kind == LR_MEDIUM_HETEROGENEOUS;
}

struct MediumInteraction {
    float t; V3 p; V3 wi; V3 sigma_s, sigma_n, sigma_t, combined_extinction; float mint; int medium;
};
LR_DEV bool mei_valid(const MediumInteraction &m) { return m.t != LR_INF; }

LR_DEV V3 medium_majorant(const DMedium &m) {
    if (m.m.kind == LR_MEDIUM_HETEROGENEOUS) return mk3(m.max_density);
    return mk3(m.m.sigma_t_const[0], m.m.sigma_t_const[1], m.m.sigma_t_const[2]) * m.m.scale;
}
LR_DEV V3 medium_albedo(const DScene &sc, const DMedium &m, V3 p) {
    if (m.m.albedo_grid >= 0) return grid_eval<false>(sc.grids[m.m.albedo_grid], p, nullptr);
    return mk3(m.m.albedo_const[0], m.m.albedo_const[1], m.m.albedo_const[2]);
}

// Medium::sample_interaction, global-majorant delta tracking step (reference draw order).
template <bool WANT_TAP>
LR_DEV MediumInteraction sample_interaction(const DScene &sc, int mid, const Ray &ray, float sample, uint32_t channel,
                                            GridTap *tap, uint32_t &n_collisions) {
    const DMedium &m = sc.media[mid]; MediumInteraction mei; mei.wi = -ray.d; mei.medium = mid;
    bool active = true; float mint = 0.f, maxt = LR_INF;
    if (m.m.kind == LR_MEDIUM_HETEROGENEOUS) {
        const DGrid &g = sc.grids[m.m.sigma_t_grid];
        bool its = aabb_intersect(g.bmin, g.bmax, ray, mint, maxt);
        its &= (isfinite(mint) || isfinite(maxt));
        active = its;
        if (!active) { mint = 0.f; maxt = LR_INF; }
    }
    mint = fmaxf(0.f, mint); maxt = fminf(ray.maxt, maxt);
    V3 ce = medium_majorant(m); float mj = idx3(ce, channel);
    float sampled_t = __fadd_rn(mint, -logf(1.f - sample) / mj);
    bool valid_mi = active && (sampled_t <= maxt);
    mei.t = valid_mi ? sampled_t : LR_INF;
    mei.p = ray_at(ray, sampled_t); mei.mint = mint;
    mei.sigma_s = mei.sigma_n = mei.sigma_t = mk3(0.f);
    if (valid_mi) {
        n_collisions++;
        V3 albedo = medium_albedo(sc, m, mei.p);
        if (m.m.kind == LR_MEDIUM_HETEROGENEOUS) {
            V3 st = grid_eval<WANT_TAP>(sc.grids[m.m.sigma_t_grid], mei.p, tap) * m.m.scale;
            mei.sigma_t = st; mei.sigma_s = st * albedo; mei.sigma_n = mk3(m.max_density) - st;
        } else {
            V3 st = mk3(m.m.sigma_t_const[0], m.m.sigma_t_const[1], m.m.sigma_t_const[2]) * m.m.scale;
            mei.sigma_t = st; mei.sigma_s = st * albedo;
        }
    }
    mei.combined_extinction = ce;
    return mei;
}

} // namespace lr
