// lr_math.cuh -- device-side fp32 vector math, PCG32/TEA streams and sampling warps.
//
// Arithmetic follows the reference's operation order where it is observable at the
// stated 1e-4 per-path tolerance (explicit fused multiply-adds where the reference
// writes dr::fmadd; IEEE division / sqrt, never the approximate intrinsics: the parity
// target is the LLVM variant).  File:line citations are relative to /root/reference.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

#define LR_DEV __device__ __forceinline__
#define LR_HD __host__ __device__ __forceinline__

namespace lr {

constexpr float kPi = 3.14159265358979323846f;
constexpr float kInvPi = 0.31830988618379067154f;
constexpr float kInvTwoPi = 0.15915494309189533577f;
constexpr float kInvFourPi = 0.07957747154594766788f;
constexpr float kInvSqrtPi = 0.56418958354775628695f;
// dr::Epsilon<float> = 2^-24; RayEpsilon = 1500 eps; ShadowEpsilon = 10 RayEpsilon
// (ext/drjit/include/drjit/array_constants.h:52-54, include/mitsuba/core/math.h:18-23)
constexpr float kEps = 5.9604644775390625e-08f;
constexpr float kRayEps = kEps * 1500.f;
constexpr float kShadowEps = kRayEps * 10.f;
#define LR_INF (__int_as_float(0x7f800000))
#define LR_LARGEST 3.402823466e+38f

struct V3 { float x, y, z; };
struct V2 { float x, y; };

LR_HD V3 mk3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
LR_HD V3 mk3(float a) { return mk3(a, a, a); }
LR_HD V2 mk2(float x, float y) { V2 r; r.x = x; r.y = y; return r; }
LR_HD V3 operator+(V3 a, V3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
LR_HD V3 operator-(V3 a, V3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
LR_HD V3 operator*(V3 a, V3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
LR_HD V3 operator/(V3 a, V3 b) { return mk3(a.x / b.x, a.y / b.y, a.z / b.z); }
LR_HD V3 operator*(V3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
LR_HD V3 operator*(float s, V3 a) { return mk3(a.x * s, a.y * s, a.z * s); }
// array / scalar is array * rcp(scalar) in Dr.Jit (ext/drjit/include/drjit/array_router.h:218-232): one correctly rounded 1 / s, then three products
LR_HD V3 operator/(V3 a, float s) { const float r = 1.f / s; return mk3(a.x * r, a.y * r, a.z * r); }
LR_HD V3 operator-(V3 a) { return mk3(-a.x, -a.y, -a.z); }
LR_HD void operator+=(V3 &a, V3 b) { a = a + b; }
LR_HD void operator*=(V3 &a, V3 b) { a = a * b; }
LR_HD void operator*=(V3 &a, float s) { a = a * s; }
LR_HD float comp(V3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

LR_DEV V3 fma3(V3 a, float b, V3 c) { return mk3(fmaf(a.x, b, c.x), fmaf(a.y, b, c.y), fmaf(a.z, b, c.z)); }
LR_DEV V3 fma3(V3 a, V3 b, V3 c) { return mk3(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y), fmaf(a.z, b.z, c.z)); }
// drjit dot(): x*y first, then fmadd chain (ext/drjit/include/drjit/array_router.h)
LR_DEV float dot(V3 a, V3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, __fmul_rn(a.x, b.x))); }
LR_DEV float abs_dot(V3 a, V3 b) { return fabsf(dot(a, b)); }
LR_DEV float sqnorm(V3 a) { return dot(a, a); }
LR_DEV float norm(V3 a) { return sqrtf(sqnorm(a)); }
LR_DEV V3 normalize(V3 a) { return a * (1.f / sqrtf(sqnorm(a))); }
// cross = fmsub(a.yzx, b.zxy, a.zxy * b.yzx) (array_router.h:671-681)
LR_DEV V3 cross(V3 a, V3 b) {
    return mk3(fmaf(a.y, b.z, -__fmul_rn(a.z, b.y)), fmaf(a.z, b.x, -__fmul_rn(a.x, b.z)), fmaf(a.x, b.y, -__fmul_rn(a.y, b.x)));
}
LR_DEV float hmax(V3 a) { return fmaxf(a.x, fmaxf(a.y, a.z)); }
LR_DEV float hmean(V3 a) { return __fmul_rn(__fadd_rn(__fadd_rn(a.x, a.y), a.z), 1.f / 3.f); }
LR_DEV bool any_nonzero(V3 a) { return a.x != 0.f || a.y != 0.f || a.z != 0.f; }
LR_DEV float sqr(float x) { return x * x; }
LR_DEV float safe_sqrt(float x) { return sqrtf(fmaxf(x, 0.f)); }
LR_DEV float mulsign(float a, float b) { return __int_as_float(__float_as_int(a) ^ (__float_as_int(b) & 0x80000000)); }
LR_DEV float mulsign_neg(float a, float b) { return __int_as_float(__float_as_int(a) ^ ((~__float_as_int(b)) & 0x80000000)); }
LR_DEV V3 mulsign(V3 a, float b) { return mk3(mulsign(a.x, b), mulsign(a.y, b), mulsign(a.z, b)); }
LR_DEV float sign_of(float x) { return copysignf(1.f, x); }
LR_DEV float lerpf(float a, float b, float t) { return fmaf(b, t, fmaf(-a, t, a)); }   // dr::lerp (array_router.h:350-352)
LR_DEV V3 exp3(V3 a) { return mk3(expf(a.x), expf(a.y), expf(a.z)); }
LR_DEV float idx3(V3 s, uint32_t ch) { return ch == 1 ? s.y : (ch == 2 ? s.z : s.x); }

struct M4 { float m[16]; };
LR_DEV V3 xform_point(const M4 &M, V3 p) {
    return mk3(fmaf(M.m[0], p.x, fmaf(M.m[1], p.y, fmaf(M.m[2], p.z, M.m[3]))),
               fmaf(M.m[4], p.x, fmaf(M.m[5], p.y, fmaf(M.m[6], p.z, M.m[7]))),
               fmaf(M.m[8], p.x, fmaf(M.m[9], p.y, fmaf(M.m[10], p.z, M.m[11]))));
}
LR_DEV V3 xform_point_proj(const M4 &M, V3 p) {
    V3 r = xform_point(M, p);
    float w = fmaf(M.m[12], p.x, fmaf(M.m[13], p.y, fmaf(M.m[14], p.z, M.m[15])));
    return r / w;
}
LR_DEV V3 xform_vector(const M4 &M, V3 v) {
    return mk3(fmaf(M.m[0], v.x, fmaf(M.m[1], v.y, __fmul_rn(M.m[2], v.z))),
               fmaf(M.m[4], v.x, fmaf(M.m[5], v.y, __fmul_rn(M.m[6], v.z))),
               fmaf(M.m[8], v.x, fmaf(M.m[9], v.y, __fmul_rn(M.m[10], v.z))));
}

// coordinate_system (include/mitsuba/core/vector.h:118-138), Duff et al. 2017
LR_DEV void coordinate_system(V3 n, V3 &s, V3 &t) {
    float sign = sign_of(n.z), a = -1.f / (sign + n.z), b = n.x * n.y * a;
    s = mk3(__fadd_rn(mulsign(__fmul_rn(sqr(n.x), a), n.z), 1.f), mulsign(b, n.z), mulsign_neg(n.x, n.z));
    t = mk3(b, fmaf(n.y, __fmul_rn(n.y, a), sign), -n.y);
}
struct Frame {
    V3 s, t, n;
    LR_DEV V3 to_local(V3 v) const { return mk3(dot(v, s), dot(v, t), dot(v, n)); }
    LR_DEV V3 to_world(V3 v) const { return fma3(n, v.z, fma3(t, v.y, s * v.x)); }
};
LR_DEV Frame frame_from_n(V3 n) { Frame f; f.n = n; coordinate_system(n, f.s, f.t); return f; }

// ---- RNG ----------------------------------------------------------------------------
// sample_tea_32 (include/mitsuba/core/random.h:76-90)
LR_HD void tea32(uint32_t v0, uint32_t v1, int rounds, uint32_t &o0, uint32_t &o1) {
    uint32_t sum = 0;
    for (int i = 0; i < rounds; ++i) {
        sum += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    o0 = v0; o1 = v1;
}
// dr::PCG32 (ext/drjit/include/drjit/random.h:135-160, 287-289) seeded as
// PCG32Sampler::seed does in the JIT variants (src/render/sampler.cpp:129-148)
// permute() (include/mitsuba/core/random.h:196-213): bit-flip permutation of [0, size), size a power of two
LR_HD uint32_t permute_pow2(uint32_t index, uint32_t size, uint32_t seed) {
    for (uint32_t bit = 1; bit < size; bit <<= 1) {
        uint32_t r0, r1; tea32(index | bit, seed, 2, r0, r1);
        if ((r0 & bit) == bit) index ^= bit;
    }
    return index;
}
// radical_inverse_2 / sobol_2, single precision (include/mitsuba/core/qmc.h:190-212, 218-251)
LR_HD uint32_t reverse_bits32(uint32_t index) {
    index = (index << 16) | (index >> 16);
    index = ((index & 0x00ff00ffu) << 8) | ((index & 0xff00ff00u) >> 8);
    index = ((index & 0x0f0f0f0fu) << 4) | ((index & 0xf0f0f0f0u) >> 4);
    index = ((index & 0x33333333u) << 2) | ((index & 0xccccccccu) >> 2);
    index = ((index & 0x55555555u) << 1) | ((index & 0xaaaaaaaau) >> 1);
    return index;
}
LR_HD uint32_t sobol_2_bits(uint32_t index, uint32_t scramble) {
    for (uint32_t v = 1u << 31; index != 0u; index >>= 1, v ^= v >> 1) if (index & 1u) scramble ^= v;
    return scramble;
}

struct Sampler {
    // independent: PCG32 (state, inc) -- inc is always odd.
    // ldsampler (src/samplers/ldsampler.cpp:107-141), marked by an even inc:
    //   state = scramble_seed | dimension_index << 32,   inc = sample_count << 1 | sample_index << 32
    uint64_t state, inc;
    LR_DEV bool is_ld() const { return (inc & 1ull) == 0ull; }
    LR_DEV uint32_t next_u32() {
        uint64_t old = state;
        state = old * 0x5851f42d4c957f2dULL + inc;
        uint32_t xs = (uint32_t) (((old >> 18u) ^ old) >> 27u);
        uint32_t rot = (uint32_t) (old >> 59u);
        return __funnelshift_r(xs, xs, rot);
    }
    LR_DEV void seed(uint32_t base_plus_seed, uint32_t lane) {
        uint32_t v0, v1; tea32(base_plus_seed, lane, 4, v0, v1);
        state = 0; inc = ((uint64_t) v1 << 1) | 1u;
        next_u32(); state += (uint64_t) v0; next_u32();
    }
    // Sampler::seed + compute_per_sequence_seed + current_sample_index (src/render/sampler.cpp:54-68, 101-121) for wavefront lane
    // `lane` with samples_per_wavefront = spp_per_pass: one sequence per pixel, sample index = lane % spp_per_pass
    LR_DEV void seed_ld(uint32_t base_seed, uint32_t seed, uint32_t lane, uint32_t spp_per_pass, uint32_t sample_count) {
        uint32_t v0, v1; tea32(base_seed, (lane / spp_per_pass) * spp_per_pass + seed, 4, v0, v1);
        state = (uint64_t) v0; inc = ((uint64_t) (lane % spp_per_pass) << 32) | ((uint64_t) sample_count << 1);
    }
    // Sampler::advance (sampler.cpp:70-73) between the passes of a multi-pass render: PCG32 streams simply continue
    LR_DEV void advance(uint32_t spp_per_pass) { if (is_ld()) { state &= 0xffffffffull; inc += (uint64_t) spp_per_pass << 32; } }
    LR_DEV uint32_t ld_index() {                     // shuffled sample index of the current dimension; bumps the dimension
        uint32_t scr = (uint32_t) state, dim = (uint32_t) (state >> 32); state += 1ull << 32;
        return permute_pow2((uint32_t) (inc >> 32), ((uint32_t) inc) >> 1, scr + dim);
    }
    LR_DEV float next_1d() {
        if (is_ld()) {
            uint32_t i = ld_index(), s0, s1; tea32((uint32_t) state, 0x48bc48ebu, 4, s0, s1);
            return __uint_as_float(((reverse_bits32(i) ^ s0) >> 9) | 0x3f800000u) - 1.f;
        }
        return __uint_as_float((next_u32() >> 9) | 0x3f800000u) - 1.f;
    }
    LR_DEV V2 next_2d() {
        if (is_ld()) {
            uint32_t i = ld_index(), sx, sy; tea32((uint32_t) state, 0x98bc51abu, 4, sx, sy);
            float x = __uint_as_float(((reverse_bits32(i) ^ sx) >> 9) | 0x3f800000u) - 1.f;
            float y = __uint2float_rn(sobol_2_bits(i, sy)) / 4294967296.f;
            return mk2(x, y);
        }
        float a = next_1d(); float b = next_1d(); return mk2(a, b);
    }
};

// ---- warps (include/mitsuba/core/warp.h) ---------------------------------------------
LR_DEV V2 square_to_uniform_disk_concentric(V2 u) {        // warp.h:54-90
    float x = fmaf(2.f, u.x, -1.f), y = fmaf(2.f, u.y, -1.f);
    bool is_zero = (x == 0.f) && (y == 0.f), q13 = fabsf(x) < fabsf(y);
    float r = q13 ? y : x, rp = q13 ? x : y;
    float phi = 0.25f * kPi * rp / r;
    if (q13) phi = 0.5f * kPi - phi;
    if (is_zero) phi = 0.f;
    float s, c; sincosf(phi, &s, &c);
    return mk2(r * c, r * s);
}
LR_DEV V3 square_to_cosine_hemisphere(V2 u) {              // warp.h:412-420
    V2 p = square_to_uniform_disk_concentric(u);
    float z = safe_sqrt(1.f - fmaf(p.y, p.y, __fmul_rn(p.x, p.x)));
    return mk3(p.x, p.y, z);
}
LR_DEV V3 square_to_uniform_sphere(V2 u) {                 // warp.h:250-255
    float z = fmaf(-2.f, u.y, 1.f), r = safe_sqrt(fmaf(-z, z, 1.f));
    float s, c; sincosf(2.f * kPi * u.x, &s, &c);
    return mk3(r * c, r * s, z);
}
LR_DEV V2 square_to_uniform_triangle(V2 u) { float t = safe_sqrt(1.f - u.x); return mk2(1.f - t, t * u.y); }

// erfinv, M. Giles' single-precision approximation, the form Dr.Jit uses
// (ext/drjit/include/drjit/math.h:1555-1578)
LR_DEV float erfinv_giles(float x) {
    float w = -logf((1.f - x) * (1.f + x));
    float w1 = w - 2.5f, w2 = sqrtf(w) - 3.f;
    float p1 = 2.81022636e-08f, p2 = -0.000200214257f;
    p1 = fmaf(p1, w1, 3.43273939e-07f); p2 = fmaf(p2, w2, 0.000100950558f);
    p1 = fmaf(p1, w1, -3.5233877e-06f); p2 = fmaf(p2, w2, 0.00134934322f);
    p1 = fmaf(p1, w1, -4.39150654e-06f); p2 = fmaf(p2, w2, -0.00367342844f);
    p1 = fmaf(p1, w1, 0.00021858087f); p2 = fmaf(p2, w2, 0.00573950773f);
    p1 = fmaf(p1, w1, -0.00125372503f); p2 = fmaf(p2, w2, -0.0076224613f);
    p1 = fmaf(p1, w1, -0.00417768164f); p2 = fmaf(p2, w2, 0.00943887047f);
    p1 = fmaf(p1, w1, 0.246640727f); p2 = fmaf(p2, w2, 1.00167406f);
    p1 = fmaf(p1, w1, 1.50140941f); p2 = fmaf(p2, w2, 2.83297682f);
    return (w < 5.f ? p1 : p2) * x;
}

LR_DEV float mis_weight(float a, float b) { a *= a; b *= b; float w = a / (a + b); return isfinite(w) ? w : 0.f; }
LR_DEV float luminance(V3 c) { return c.x * 0.212671f + c.y * 0.715160f + c.z * 0.072169f; }

} // namespace lr
