/*
 * TEST INFRASTRUCTURE ONLY (see orc_math.h).  Reconstruction filters, the
 * ImageBlock splat (JIT-variant semantics) and HDRFilm::develop, restated from
 * the reference.
 */
#pragma once
#include "orc_scene.h"
#include <atomic>

namespace orc {

/* GaussianFilter ctor, non-CUDA branch (src/rfilters/gaussian.cpp:60-96):
   Remez fit of exp(-x/2) scaled by stddev, shifted to reach zero at the radius */
inline void gaussian_coeffs(float stddev, float radius, float out[10]) {
    const double coeff[10] = {9.992604880e-1, -4.977025247e-1, 1.222248550e-1, -1.932406282e-2, 2.136713061e-3,
                              -1.679873860e-4, 9.202145248e-6, -3.329417433e-7, 7.128382794e-9, -6.821193280e-11};
    double scale = 1;
    for (int i = 0; i < 10; ++i) { out[i] = (float) (coeff[i] * scale); scale /= ((double) stddev) * ((double) stddev); }
    /* dr::detail::estrin_impl(x, coeff): Estrin's scheme, fp32 */
    auto estrin = [](float x, const float *c) {
        float x2 = x * x, x4 = x2 * x2, x8 = x4 * x4;
        float p01 = fmaf(c[1], x, c[0]), p23 = fmaf(c[3], x, c[2]), p45 = fmaf(c[5], x, c[4]), p67 = fmaf(c[7], x, c[6]), p89 = fmaf(c[9], x, c[8]);
        float p0123 = fmaf(p23, x2, p01), p4567 = fmaf(p67, x2, p45);
        float p0_7 = fmaf(p4567, x4, p0123);
        return fmaf(p89, x8, p0_7);
    };
    out[0] -= estrin(radius * radius, out);
}
inline float estrin10(float x, const float *c) {
    float x2 = x * x, x4 = x2 * x2, x8 = x4 * x4;
    float p01 = fmaf(c[1], x, c[0]), p23 = fmaf(c[3], x, c[2]), p45 = fmaf(c[5], x, c[4]), p67 = fmaf(c[7], x, c[6]), p89 = fmaf(c[9], x, c[8]);
    float p0123 = fmaf(p23, x2, p01), p4567 = fmaf(p67, x2, p45);
    float p0_7 = fmaf(p4567, x4, p0123);
    return fmaf(p89, x8, p0_7);
}

/* rfilter->eval(x): analytic form used by the JIT variants
   (gaussian.cpp:92-96 LLVM branch, box.cpp:33-41, tent.cpp:45-53) */
inline float rfilter_eval(const Scene &sc, float x) {
    const lr_sensor &s = sc.sensor;
    switch (s.rfilter) {
        case LR_RFILTER_BOX: return (x >= -.5f && x < .5f) ? 1.f : 0.f;   /* box.cpp: x in [-r, r) */
        case LR_RFILTER_TENT: return std::max(0.f, 1.f - fabsf(x * (1.f / s.rfilter_radius)));
        default: return std::max(estrin10(x * x, sc.gauss_coeff), 0.f);
    }
}

/* Film accumulator: [H][W][C] doubles with lock-free adds so host threads can
   share it (the reference uses fp32 atomics in unspecified order; the oracle
   accumulates in double to be order-independent). */
struct Film {
    uint32_t W, H, C; std::vector<std::atomic<double>> acc;
    Film(uint32_t w, uint32_t h, uint32_t c) : W(w), H(h), C(c), acc((size_t) w * h * c) { clear(); }
    void clear() { for (auto &a : acc) a.store(0.0, std::memory_order_relaxed); }
    void add(size_t i, double v) {
        double old = acc[i].load(std::memory_order_relaxed);
        while (!acc[i].compare_exchange_weak(old, old + v, std::memory_order_relaxed)) {}
    }
};

/* filter footprint of one sample: pixel window + separable weights.
   ImageBlock::put (src/render/imageblock.cpp:174-560), JIT semantics:
   box fast path (:206-235); coalesced window when spp >= 4 (:431-455);
   else the clipped radius window (:262-275). `count` <= 5 on this path. */
struct Footprint { int x0, y0, nx, ny; float wx[16], wy[16]; };    /* radius <= 7.5 px */
inline bool film_footprint(const Scene &sc, V2 pos, bool coalesce, Footprint &fp) {
    const lr_sensor &s = sc.sensor; int W = (int) s.film_width, H = (int) s.film_height;
    if (s.rfilter == LR_RFILTER_BOX) {
        /* pos is the integer pixel (jitter ignored); border 0, offset 0 */
        int x = (int) floorf(pos.x), y = (int) floorf(pos.y);
        if (x < 0 || y < 0 || x >= W || y >= H) return false;
        fp.x0 = x; fp.y0 = y; fp.nx = fp.ny = 1; fp.wx[0] = fp.wy[0] = 1.f; return true;
    }
    float radius = s.rfilter_radius;
    if (coalesce) {
        int n = (int) ceilf(radius - .5f), count = 2 * n + 1;
        int px = (int) floorf(pos.x) - n, py = (int) floorf(pos.y) - n;
        float rx = ((float) px + .5f) - pos.x, ry = ((float) py + .5f) - pos.y;
        fp.x0 = px; fp.y0 = py; fp.nx = fp.ny = count;
        for (int i = 0; i < count; ++i) { fp.wx[i] = rfilter_eval(sc, rx); fp.wy[i] = rfilter_eval(sc, ry); rx += 1.f; ry += 1.f; }
        return true;
    } else {
        float pfx = pos.x - .5f, pfy = pos.y - .5f;
        int x0 = std::max((int) ceilf(pfx - radius), 0), y0 = std::max((int) ceilf(pfy - radius), 0);
        int x1 = std::min((int) floorf(pfx + radius), W - 1), y1 = std::min((int) floorf(pfy + radius), H - 1);
        if (x0 > x1 || y0 > y1) return false;
        fp.x0 = x0; fp.y0 = y0; fp.nx = x1 - x0 + 1; fp.ny = y1 - y0 + 1;
        float rx = (float) x0 - pfx, ry = (float) y0 - pfy;
        for (int i = 0; i < fp.nx; ++i) { fp.wx[i] = rfilter_eval(sc, rx); rx += 1.f; }
        for (int i = 0; i < fp.ny; ++i) { fp.wy[i] = rfilter_eval(sc, ry); ry += 1.f; }
        return true;
    }
}
inline void film_put(const Scene &sc, Film &film, V2 pos, const float *values, bool coalesce) {
    Footprint fp; if (!film_footprint(sc, pos, coalesce, fp)) return;
    for (int j = 0; j < fp.ny; ++j) {
        int y = fp.y0 + j; if (y < 0 || y >= (int) film.H) continue;
        for (int i = 0; i < fp.nx; ++i) {
            int x = fp.x0 + i; if (x < 0 || x >= (int) film.W) continue;
            float w = fp.wy[j] * fp.wx[i];
            size_t idx = ((size_t) y * film.W + x) * film.C;
            for (uint32_t k = 0; k < film.C; ++k) film.add(idx + k, (double) (values[k] * w));
        }
    }
}
/* HDRFilm::develop (src/films/hdrfilm.cpp:306-407): value / (W == 0 ? 1 : W) */
inline void film_develop(const Film &film, float *out) {
    uint32_t oc = film.C - 1;
    for (size_t p = 0; p < (size_t) film.W * film.H; ++p) {
        float w = (float) film.acc[p * film.C + film.C - 1].load(std::memory_order_relaxed);
        float d = w == 0.f ? 1.f : w;
        for (uint32_t k = 0; k < oc; ++k) out[p * oc + k] = (float) film.acc[p * film.C + k].load(std::memory_order_relaxed) / d;
    }
}

} // namespace orc
