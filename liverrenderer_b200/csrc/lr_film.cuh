// lr_film.cuh -- reconstruction-filter splat into the RGB[A]W accumulator and develop (device code).
// ImageBlock::put, JIT-variant semantics (src/render/imageblock.cpp:174-560: box fast path :206-235,
// clipped window :262-429 when spp < 4, coalesced (2n+1)^2 window :431-560 when spp >= 4);
// analytic filters (src/rfilters/gaussian.cpp:92-96 LLVM polynomial form, box.cpp:33-41, tent.cpp:45-53);
// HDRFilm::develop weight division (src/films/hdrfilm.cpp:306-407).
#pragma once
#include "lr_scene.cuh"

namespace lr {

LR_DEV float estrin10(float x, const float *c) {                 // dr::detail::estrin_impl, n = 10
    float x2 = x * x, x4 = x2 * x2, x8 = x4 * x4;
    float p01 = fmaf(c[1], x, c[0]), p23 = fmaf(c[3], x, c[2]), p45 = fmaf(c[5], x, c[4]), p67 = fmaf(c[7], x, c[6]), p89 = fmaf(c[9], x, c[8]);
    float p0123 = fmaf(p23, x2, p01), p4567 = fmaf(p67, x2, p45);
    return fmaf(p89, x8, fmaf(p4567, x4, p0123));
}
LR_DEV float rfilter_eval(const DScene &sc, float x) {
    switch (sc.sensor.rfilter) {
        case LR_RFILTER_BOX: return (x >= -.5f && x < .5f) ? 1.f : 0.f;
        case LR_RFILTER_TENT: return fmaxf(0.f, 1.f - fabsf(x * (1.f / sc.sensor.rfilter_radius)));
        default: return fmaxf(estrin10(__fmul_rn(x, x), sc.gauss_coeff), 0.f);
    }
}

struct Footprint { int x0, y0, nx, ny; float wx[6], wy[6]; };

LR_DEV bool film_footprint(const DScene &sc, V2 pos, bool coalesce, Footprint &fp) {
    const int W = (int) sc.sensor.film_width, H = (int) sc.sensor.film_height;
    if (sc.sensor.rfilter == LR_RFILTER_BOX) {
        int x = (int) floorf(pos.x), y = (int) floorf(pos.y);
        if (x < 0 || y < 0 || x >= W || y >= H) return false;
        fp.x0 = x; fp.y0 = y; fp.nx = fp.ny = 1; fp.wx[0] = fp.wy[0] = 1.f; return true;
    }
    float radius = sc.sensor.rfilter_radius;
    if (coalesce) {
        int n = (int) ceilf(radius - .5f), count = min(2 * n + 1, 6);
        int px = (int) floorf(pos.x) - n, py = (int) floorf(pos.y) - n;
        float rx = __fsub_rn(__fadd_rn((float) px, .5f), pos.x), ry = __fsub_rn(__fadd_rn((float) py, .5f), pos.y);
        fp.x0 = px; fp.y0 = py; fp.nx = fp.ny = count;
#pragma unroll
        for (int i = 0; i < 6; ++i) if (i < count) { fp.wx[i] = rfilter_eval(sc, rx); fp.wy[i] = rfilter_eval(sc, ry); rx += 1.f; ry += 1.f; }
        return true;
    }
    float pfx = pos.x - .5f, pfy = pos.y - .5f;
    int x0 = max((int) ceilf(pfx - radius), 0), y0 = max((int) ceilf(pfy - radius), 0);
    int x1 = min((int) floorf(pfx + radius), W - 1), y1 = min((int) floorf(pfy + radius), H - 1);
    if (x0 > x1 || y0 > y1) return false;
    fp.x0 = x0; fp.y0 = y0; fp.nx = min(x1 - x0 + 1, 6); fp.ny = min(y1 - y0 + 1, 6);
    float rx = (float) x0 - pfx, ry = (float) y0 - pfy;
#pragma unroll
    for (int i = 0; i < 6; ++i) { if (i < fp.nx) { fp.wx[i] = rfilter_eval(sc, rx); rx += 1.f; } if (i < fp.ny) { fp.wy[i] = rfilter_eval(sc, ry); ry += 1.f; } }
    return true;
}

// Filters wider than the 6-tap footprint (radius > 2.5 px, e.g. a gaussian with stddev 1): the same two windows of ImageBlock::put with the
// separable weights evaluated inside the loops instead of being kept in registers -- the narrow filters of the configured scenes keep their code.
template <int C>
LR_DEV void film_put_wide(const DScene &sc, float *film, V2 pos, const float *values, bool coalesce) {
    const int W = (int) sc.sensor.film_width, H = (int) sc.sensor.film_height; const float radius = sc.sensor.rfilter_radius;
    int x0, y0, nx, ny; float rx0, ry0;
    if (coalesce) {                                                // imageblock.cpp:431-455
        const int n = (int) ceilf(radius - .5f); nx = ny = 2 * n + 1;
        x0 = (int) floorf(pos.x) - n; y0 = (int) floorf(pos.y) - n;
        rx0 = __fsub_rn(__fadd_rn((float) x0, .5f), pos.x); ry0 = __fsub_rn(__fadd_rn((float) y0, .5f), pos.y);
    } else {                                                       // :262-275
        const float pfx = pos.x - .5f, pfy = pos.y - .5f;
        x0 = max((int) ceilf(pfx - radius), 0); y0 = max((int) ceilf(pfy - radius), 0);
        const int x1 = min((int) floorf(pfx + radius), W - 1), y1 = min((int) floorf(pfy + radius), H - 1);
        if (x0 > x1 || y0 > y1) return;
        nx = x1 - x0 + 1; ny = y1 - y0 + 1; rx0 = (float) x0 - pfx; ry0 = (float) y0 - pfy;
    }
    float ry = ry0;
    for (int j = 0; j < ny; ++j, ry += 1.f) {
        const int y = y0 + j; if (y < 0 || y >= H) continue;
        const float wy = rfilter_eval(sc, ry); float rx = rx0;
        for (int i = 0; i < nx; ++i, rx += 1.f) {
            const int x = x0 + i; if (x < 0 || x >= W) continue;
            const float w = __fmul_rn(wy, rfilter_eval(sc, rx));
            float *dst = film + ((size_t) y * W + x) * C;
#pragma unroll
            for (int k = 0; k < C; ++k) atomicAdd(dst + k, __fmul_rn(values[k], w));
        }
    }
}

// ImageBlock::accum -> fire-and-forget fp32 reductions into L2 (RED.E.ADD.F32)
template <int C>
LR_DEV void film_put(const DScene &sc, float *film, V2 pos, const float *values, bool coalesce) {
    if (sc.sensor.rfilter != LR_RFILTER_BOX && sc.sensor.rfilter_radius > 2.5f) { film_put_wide<C>(sc, film, pos, values, coalesce); return; }
    Footprint fp; if (!film_footprint(sc, pos, coalesce, fp)) return;
    const int W = (int) sc.sensor.film_width, H = (int) sc.sensor.film_height;
    for (int j = 0; j < fp.ny; ++j) {
        int y = fp.y0 + j; if (y < 0 || y >= H) continue;
        for (int i = 0; i < fp.nx; ++i) {
            int x = fp.x0 + i; if (x < 0 || x >= W) continue;
            float w = __fmul_rn(fp.wy[j], fp.wx[i]);
            float *dst = film + ((size_t) y * W + x) * C;
#pragma unroll
            for (int k = 0; k < C; ++k) atomicAdd(dst + k, __fmul_rn(values[k], w));
        }
    }
}

} // namespace lr
