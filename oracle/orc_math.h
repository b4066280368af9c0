/*
 * TEST INFRASTRUCTURE ONLY -- CPU oracle for the path/volpath/prbvolpath hot path.
 * Nothing under oracle/ is imported, linked or executed by the product
 * (liverrenderer_b200/); only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, as the checker.
 *
 * orc_math.h: scalar fp32 vector math, RNG and warps, restated from the
 * reference (file:line cited per function; paths relative to /root/reference).
 * Compiled with -ffp-contract=off; fused multiply-adds appear only where the
 * reference writes dr::fmadd / fmsub / fnmadd explicitly.
 */
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <algorithm>
#include <limits>

namespace orc {

static constexpr float Pi        = 3.14159265358979323846f;
static constexpr float InvPi     = 0.31830988618379067154f;
static constexpr float InvTwoPi  = 0.15915494309189533577f;
static constexpr float InvFourPi = 0.07957747154594766788f;
static constexpr float InvSqrtPi = 0.56418958354775628695f;
static constexpr float Inf       = std::numeric_limits<float>::infinity();
static constexpr float Largest   = std::numeric_limits<float>::max();
/* dr::Epsilon<float> = 2^-24 (ext/drjit/include/drjit/array_constants.h:52-54);
   RayEpsilon = Epsilon*1500, ShadowEpsilon = RayEpsilon*10
   (include/mitsuba/core/math.h:18-23) */
static constexpr float Epsilon       = 5.9604644775390625e-08f;
static constexpr float RayEpsilon    = Epsilon * 1500.f;
static constexpr float ShadowEpsilon = RayEpsilon * 10.f;

struct V2 { float x, y; };
struct V3 {
    float x, y, z;
    V3() : x(0), y(0), z(0) {}
    V3(float a) : x(a), y(a), z(a) {}
    V3(float a, float b, float c) : x(a), y(b), z(c) {}
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
    float &operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
};
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline V3 operator/(V3 a, V3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator*(float s, V3 a) { return {a.x * s, a.y * s, a.z * s}; }
/* array / scalar is array * rcp(scalar) in Dr.Jit (ext/drjit/include/drjit/array_router.h:218-232), one rounding of 1 / s shared by the lanes */
inline V3 operator/(V3 a, float s) { float r = 1.f / s; return {a.x * r, a.y * r, a.z * r}; }
inline V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
inline V3 &operator+=(V3 &a, V3 b) { a = a + b; return a; }
inline V3 &operator*=(V3 &a, V3 b) { a = a * b; return a; }
inline V3 &operator*=(V3 &a, float s) { a = a * s; return a; }
inline V3 fma3(V3 a, float b, V3 c) { return {fmaf(a.x, b, c.x), fmaf(a.y, b, c.y), fmaf(a.z, b, c.z)}; }
inline V3 fma3(V3 a, V3 b, V3 c) { return {fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y), fmaf(a.z, b.z, c.z)}; }
/* drjit dot(): fmadd chain, x*y first (ext/drjit/include/drjit/array_router.h, horizontal ops) */
inline float dot(V3 a, V3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
inline float abs_dot(V3 a, V3 b) { return fabsf(dot(a, b)); }
inline float sqnorm(V3 a) { return dot(a, a); }
inline float norm(V3 a) { return sqrtf(sqnorm(a)); }
inline V3 normalize(V3 a) { return a * (1.f / sqrtf(sqnorm(a))); }
inline V3 cross(V3 a, V3 b) {
    return {fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x))};
}
inline float hmax(V3 a) { return std::max(a.x, std::max(a.y, a.z)); }
inline float hmean(V3 a) { return (a.x + a.y + a.z) * (1.f / 3.f); }
inline bool any_nonzero(V3 a) { return a.x != 0.f || a.y != 0.f || a.z != 0.f; }
inline float sqr(float x) { return x * x; }
inline float safe_sqrt(float x) { return sqrtf(std::max(x, 0.f)); }
inline float mulsign(float a, float b) { return std::signbit(b) ? -a : a; }
inline float mulsign_neg(float a, float b) { return std::signbit(b) ? a : -a; }
inline V3 mulsign(V3 a, float b) { return {mulsign(a.x, b), mulsign(a.y, b), mulsign(a.z, b)}; }
inline float sign_of(float x) { return std::signbit(x) ? -1.f : 1.f; }   /* dr::sign: copysign(1, x) */
inline float lerp(float a, float b, float t) { return fmaf(b, t, fmaf(-a, t, a)); }

/* 4x4 row-major affine / projective transforms */
struct M4 { float m[16]; };
inline V3 xform_point_affine(const M4 &M, V3 p) {
    return {fmaf(M.m[0], p.x, fmaf(M.m[1], p.y, fmaf(M.m[2], p.z, M.m[3]))),
            fmaf(M.m[4], p.x, fmaf(M.m[5], p.y, fmaf(M.m[6], p.z, M.m[7]))),
            fmaf(M.m[8], p.x, fmaf(M.m[9], p.y, fmaf(M.m[10], p.z, M.m[11])))};
}
inline V3 xform_point_proj(const M4 &M, V3 p) {
    V3 r = xform_point_affine(M, p);
    float w = fmaf(M.m[12], p.x, fmaf(M.m[13], p.y, fmaf(M.m[14], p.z, M.m[15])));
    return r / w;
}
inline V3 xform_vector(const M4 &M, V3 v) {
    return {fmaf(M.m[0], v.x, fmaf(M.m[1], v.y, M.m[2] * v.z)),
            fmaf(M.m[4], v.x, fmaf(M.m[5], v.y, M.m[6] * v.z)),
            fmaf(M.m[8], v.x, fmaf(M.m[9], v.y, M.m[10] * v.z))};
}

/* coordinate_system (include/mitsuba/core/vector.h:118-138), Duff et al. 2017 */
inline void coordinate_system(V3 n, V3 &s, V3 &t) {
    float sign = sign_of(n.z), a = -1.f / (sign + n.z), b = n.x * n.y * a;
    s = V3(mulsign(sqr(n.x) * a, n.z) + 1.f, mulsign(b, n.z), mulsign_neg(n.x, n.z));
    t = V3(b, fmaf(n.y, n.y * a, sign), -n.y);
}

/* Frame3f (include/mitsuba/core/frame.h) */
struct Frame {
    V3 s, t, n;
    Frame() {}
    explicit Frame(V3 n_) : n(n_) { coordinate_system(n_, s, t); }
    V3 to_local(V3 v) const { return {dot(v, s), dot(v, t), dot(v, n)}; }
    V3 to_world(V3 v) const { return fma3(n, v.z, fma3(t, v.y, s * v.x)); }
};

/* ---- RNG ------------------------------------------------------------------ */
/* sample_tea_32 (include/mitsuba/core/random.h:76-90) */
inline void tea32(uint32_t v0, uint32_t v1, int rounds, uint32_t &o0, uint32_t &o1) {
    uint32_t sum = 0;
    for (int i = 0; i < rounds; ++i) {
        sum += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    o0 = v0; o1 = v1;
}
/* sample_tea_float32 (random.h:118-121): (v1 >> 9) | 0x3f800000 - 1 */
inline float tea_float32(uint32_t v0, uint32_t v1, int rounds) {
    uint32_t a, b; tea32(v0, v1, rounds, a, b);
    uint32_t bits = (b >> 9) | 0x3f800000u; float f; memcpy(&f, &bits, 4); return f - 1.f;
}

/* dr::PCG32 (ext/drjit/include/drjit/random.h:103-300) */
struct PCG32 {
    static constexpr uint64_t MULT = 0x5851f42d4c957f2dULL;
    static constexpr uint64_t DEFAULT_STATE  = 0x853c49e6748fea9bULL;
    static constexpr uint64_t DEFAULT_STREAM = 0xda3e39cb94b95bdbULL;
    uint64_t state, inc;
    void seed(uint64_t initstate = DEFAULT_STATE, uint64_t initseq = DEFAULT_STREAM) {
        state = 0; inc = (initseq << 1) | 1u;
        next_u32(); state += initstate; next_u32();
    }
    uint32_t next_u32() {
        uint64_t old = state;
        state = old * MULT + inc;
        uint32_t xs = (uint32_t) (((old >> 18u) ^ old) >> 27u);
        uint32_t rot = (uint32_t) (old >> 59u);
        return (xs >> rot) | (xs << ((-rot) & 31));
    }
    float next_f32() {
        uint32_t bits = (next_u32() >> 9) | 0x3f800000u; float f; memcpy(&f, &bits, 4); return f - 1.f;
    }
};

/* independent sampler in the JIT variants: PCG32Sampler::seed
   (src/render/sampler.cpp:129-148) + IndependentSampler::next_1d/2d
   (src/samplers/independent.cpp:88-97) */
/* permute() for power-of-two sizes (include/mitsuba/core/random.h:196-213) */
inline uint32_t permute_pow2(uint32_t index, uint32_t size, uint32_t seed) {
    uint32_t n = 0; while ((1u << n) < size) ++n;
    for (uint32_t level = 0; level < n; ++level) {
        uint32_t bit = 1u << level, r0, r1; tea32(index | bit, seed, 2, r0, r1);
        if ((r0 & bit) == bit) index = index ^ bit;
    }
    return index;
}
/* radical_inverse_2 / sobol_2, single precision (include/mitsuba/core/qmc.h:190-212, 218-251) */
inline float radical_inverse_2(uint32_t index, uint32_t scramble) {
    index = (index << 16) | (index >> 16);
    index = ((index & 0x00ff00ffu) << 8) | ((index & 0xff00ff00u) >> 8);
    index = ((index & 0x0f0f0f0fu) << 4) | ((index & 0xf0f0f0f0u) >> 4);
    index = ((index & 0x33333333u) << 2) | ((index & 0xccccccccu) >> 2);
    index = ((index & 0x55555555u) << 1) | ((index & 0xaaaaaaaau) >> 1);
    uint32_t bits = ((index ^ scramble) >> 9) | 0x3f800000u; float f; memcpy(&f, &bits, 4); return f - 1.f;
}
inline float sobol_2(uint32_t index, uint32_t scramble) {
    uint32_t v = 1u << 31;
    while (index != 0u) { if ((index & 1u) == 1u) scramble ^= v; index >>= 1; v ^= v >> 1; }
    return (float) scramble / (float) (1ull << 32);
}

/* The sensor's sampler: `independent` (PCG32Sampler, src/render/sampler.cpp:129-148) or `ldsampler`
   (src/samplers/ldsampler.cpp:107-141 on top of Sampler::seed / advance / compute_per_sequence_seed /
   current_sample_index, src/render/sampler.cpp:54-121). */
struct Sampler {
    PCG32 rng;
    bool ld = false; uint32_t scramble_seed = 0, dimension_index = 0, sample_index = 0, sample_count = 0;
    void seed(uint32_t base_plus_seed, uint32_t lane) {
        uint32_t v0, v1; tea32(base_plus_seed, lane, 4, v0, v1);
        rng.seed((uint64_t) v0, (uint64_t) v1); ld = false;
    }
    /* wavefront lane `lane`, samples_per_wavefront = spp: sequence = pixel, sample index = lane % spp */
    void seed_ld(uint32_t base_seed, uint32_t seed, uint32_t lane, uint32_t spp, uint32_t count) {
        uint32_t v0, v1; tea32(base_seed, spp * (lane / spp) + seed, 4, v0, v1);
        ld = true; scramble_seed = v0; dimension_index = 0; sample_index = lane % spp; sample_count = count;
    }
    void advance(uint32_t samples_per_wavefront) { dimension_index = 0; sample_index += samples_per_wavefront; }
    float next_1d() {
        if (!ld) return rng.next_f32();
        uint32_t perm_seed = scramble_seed + dimension_index++;
        uint32_t i = permute_pow2(sample_index, sample_count, perm_seed);
        uint32_t s0, s1; tea32(scramble_seed, 0x48bc48ebu, 4, s0, s1);
        return radical_inverse_2(i, s0);
    }
    V2 next_2d() {
        if (!ld) { float a = rng.next_f32(); float b = rng.next_f32(); return {a, b}; }
        uint32_t perm_seed = scramble_seed + dimension_index++;
        uint32_t i = permute_pow2(sample_index, sample_count, perm_seed);
        uint32_t sx, sy; tea32(scramble_seed, 0x98bc51abu, 4, sx, sy);
        return {radical_inverse_2(i, sx), sobol_2(i, sy)};
    }
};

/* ---- warps (include/mitsuba/core/warp.h) ---------------------------------- */
inline void sincos_(float x, float &s, float &c) { s = sinf(x); c = cosf(x); }
/* warp.h:54-90 */
inline V2 square_to_uniform_disk_concentric(V2 u) {
    float x = fmaf(2.f, u.x, -1.f), y = fmaf(2.f, u.y, -1.f);
    bool is_zero = (x == 0.f) && (y == 0.f), q13 = fabsf(x) < fabsf(y);
    float r = q13 ? y : x, rp = q13 ? x : y;
    float phi = 0.25f * Pi * rp / r;
    if (q13) phi = 0.5f * Pi - phi;
    if (is_zero) phi = 0.f;
    float s, c; sincos_(phi, s, c);
    return {r * c, r * s};
}
/* warp.h:412-420 */
inline V3 square_to_cosine_hemisphere(V2 u) {
    V2 p = square_to_uniform_disk_concentric(u);
    float z = safe_sqrt(1.f - fmaf(p.y, p.y, p.x * p.x));
    return {p.x, p.y, z};
}
inline float square_to_cosine_hemisphere_pdf(V3 v) { return InvPi * v.z; }
/* warp.h:250-255 */
inline V3 square_to_uniform_sphere(V2 u) {
    float z = fmaf(-2.f, u.y, 1.f), r = safe_sqrt(fmaf(-z, z, 1.f));
    float s, c; sincos_(2.f * Pi * u.x, s, c);
    return {r * c, r * s, z};
}
/* warp.h square_to_uniform_triangle */
inline V2 square_to_uniform_triangle(V2 u) {
    float t = safe_sqrt(1.f - u.x);
    return {1.f - t, t * u.y};
}

/* erfinv: "Approximating the erfinv function", M. Giles -- the form Dr.Jit
   uses (ext/drjit/include/drjit/math.h:1555-1578); estrin evaluation order is
   replaced by Horner (difference below 1 ulp of the result's magnitude). */
inline float erfinv_(float x) {
    float w = -logf((1.f - x) * (1.f + x));
    float w1 = w - 2.5f, w2 = sqrtf(w) - 3.f;
    const float c1[9] = {1.50140941f, 0.246640727f, -0.00417768164f, -0.00125372503f, 0.00021858087f,
                         -4.39150654e-06f, -3.5233877e-06f, 3.43273939e-07f, 2.81022636e-08f};
    const float c2[9] = {2.83297682f, 1.00167406f, 0.00943887047f, -0.0076224613f, 0.00573950773f,
                         -0.00367342844f, 0.00134934322f, 0.000100950558f, -0.000200214257f};
    float p1 = c1[8], p2 = c2[8];
    for (int i = 7; i >= 0; --i) { p1 = fmaf(p1, w1, c1[i]); p2 = fmaf(p2, w2, c2[i]); }
    return (w < 5.f ? p1 : p2) * x;
}

} // namespace orc
