// ref_kat.cpp -- the reference's OWN header-only code for the hot path's arithmetic, compiled for Float = float, Spectrum = Color3f
// straight from /root/reference (nothing is copied into this repository): drjit/random.h (PCG32), mitsuba/core/random.h (TEA),
// core/warp.h, render/fresnel.h, render/microfacet.h, src/bsdfs/principledhelpers.h, core/vector.h, drjit/math.h.
// Built by oracle/ref_kat/Makefile into oracle/_ref/libref_kat.so; tests/test_oracle_ref.py asserts that the oracle's hand
// restatement (orc_kat_fn) returns the same bits.  TEST INFRASTRUCTURE ONLY.
#include <drjit/random.h>
#include <drjit/math.h>
#include <mitsuba/core/random.h>
#include <mitsuba/core/vector.h>
#include <mitsuba/core/warp.h>
#include <mitsuba/render/fresnel.h>
#include <mitsuba/render/microfacet.h>
#include <src/bsdfs/principledhelpers.h>
#include <cstring>
#include "../kat_fn.h"

using namespace mitsuba;
// Float = a one-lane packet, not plain float: with a scalar Float Dr.Jit takes shortcuts that the JIT variants do not (squared_norm /
// dot as sum(v * v) instead of an fmadd chain, libm instead of its own polynomials), and the path's target is llvm_ad_rgb.  A static
// array of one lane has the array depth of the JIT types, so every template takes the same branches as under the JIT.
using Flt = dr::Packet<float, 1>;
using Spec = Color<Flt, 3>;
using Vector3f = Vector<Flt, 3>;
using Normal3f = Normal<Flt, 3>;
using Point2f = Point<Flt, 2>;
using U32 = dr::Packet<uint32_t, 1>;
using MF = MicrofacetDistribution<Flt, Spec>;

static uint32_t bits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static float fbits(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static Vector3f v3(const float *p) { return Vector3f(Flt(p[0]), Flt(p[1]), Flt(p[2])); }
static void put3(float *o, const Vector3f &v) { o[0] = v.x()[0]; o[1] = v.y()[0]; o[2] = v.z()[0]; }
static float f1(const Flt &v) { return v[0]; }
static MF mf(const float *in) { return MF(in[0] == 0.f ? MicrofacetType::Beckmann : MicrofacetType::GGX, Flt(in[1]), Flt(in[2]), in[3] != 0.f); }

extern "C" int ref_kat_fn(int fn, const float *in, float *out) {
    using Mask = dr::mask_t<Flt>;
    switch (fn) {
        case KAT_TEA32: { auto r = sample_tea_32<U32>(U32(bits(in[0])), U32(bits(in[1])), (int) bits(in[2])); out[0] = fbits(r.first[0]); out[1] = fbits(r.second[0]); return 2; }
        case KAT_TEA_FLOAT32: out[0] = f1(sample_tea_float32<U32>(U32(bits(in[0])), U32(bits(in[1])), (int) bits(in[2]))); return 1;
        case KAT_PCG32: {
            using U64 = dr::Packet<uint64_t, 1>;
            dr::PCG32<U32> rng; rng.seed(U64(((uint64_t) bits(in[1]) << 32) | bits(in[0])), U64(((uint64_t) bits(in[3]) << 32) | bits(in[2])));
            for (int i = 0; i < 4; ++i) out[i] = rng.next_float32()[0];
            for (int i = 0; i < 4; ++i) out[4 + i] = fbits(rng.next_uint32()[0]);
            return 8; }
        case KAT_WARP_DISK: { auto p = warp::square_to_uniform_disk_concentric(Point2f(Flt(in[0]), Flt(in[1]))); out[0] = p.x()[0]; out[1] = p.y()[0]; return 2; }
        case KAT_WARP_COS_HEMI: { auto v = warp::square_to_cosine_hemisphere(Point2f(Flt(in[0]), Flt(in[1]))); put3(out, v); out[3] = f1(warp::square_to_cosine_hemisphere_pdf(v)); return 4; }
        case KAT_WARP_SPHERE: { auto v = warp::square_to_uniform_sphere(Point2f(Flt(in[0]), Flt(in[1]))); put3(out, v); return 3; }
        case KAT_WARP_TRIANGLE: { auto p = warp::square_to_uniform_triangle(Point2f(Flt(in[0]), Flt(in[1]))); out[0] = p.x()[0]; out[1] = p.y()[0]; return 2; }
        case KAT_FRESNEL: { auto [F, ct, eit, eti] = fresnel(Flt(in[0]), Flt(in[1])); out[0] = f1(F); out[1] = f1(ct); out[2] = f1(eit); out[3] = f1(eti); return 4; }
        case KAT_REFLECT: put3(out, reflect(v3(in), Normal3f(v3(in + 3)))); return 3;
        case KAT_REFRACT: put3(out, refract(v3(in), Normal3f(v3(in + 3)), Flt(in[6]), Flt(in[7]))); return 3;
        case KAT_MF_EVAL: out[0] = f1(mf(in).eval(v3(in + 4))); return 1;
        case KAT_MF_PDF: out[0] = f1(mf(in).pdf(v3(in + 4), v3(in + 7))); return 1;
        case KAT_MF_SAMPLE: { auto [m, pdf] = mf(in).sample(v3(in + 4), Point2f(Flt(in[7]), Flt(in[8]))); put3(out, m); out[3] = f1(pdf); return 4; }
        case KAT_MF_G1: out[0] = f1(mf(in).smith_g1(v3(in + 4), v3(in + 7))); return 1;
        case KAT_MF_G: out[0] = f1(mf(in).G(v3(in + 4), v3(in + 7), v3(in + 10))); return 1;
        case KAT_GTR1: { GTR1Isotropic<Flt, Spec> g{Flt(in[0])}; out[0] = f1(g.eval(v3(in + 1))); out[1] = f1(g.pdf(v3(in + 1))); put3(out + 2, g.sample(Point2f(Flt(in[4]), Flt(in[5])))); return 5; }
        case KAT_SMITH_GGX1: out[0] = f1(smith_ggx1<Flt>(v3(in), v3(in + 3), Flt(in[6]))); out[1] = f1(clearcoat_G<Flt>(v3(in), v3(in + 7), v3(in + 3), Flt(in[6]))); return 2;
        case KAT_SCHLICK: out[0] = f1(schlick_weight<Flt>(Flt(in[0]))); out[1] = f1(calc_schlick<Flt, Flt>(Flt(in[1]), Flt(in[0]), Flt(in[2]))); out[2] = f1(schlick_R0_eta<Flt>(Flt(in[2]))); return 3;
        case KAT_MAC_MIC: out[0] = dr::all(mac_mic_compatibility<Flt>(v3(in), v3(in + 3), v3(in + 6), Flt(in[9]), true)) ? 1.f : 0.f;
                          out[1] = dr::all(mac_mic_compatibility<Flt>(v3(in), v3(in + 3), v3(in + 6), Flt(in[9]), false)) ? 1.f : 0.f; return 2;
        case KAT_PRINCIPLED_FRESNEL: {
            Spec r = principled_fresnel<Flt, Spec>(Flt(in[0]), Flt(in[1]), Flt(in[2]), Spec(Flt(in[3]), Flt(in[4]), Flt(in[5])), Flt(in[6]), Flt(in[7]), Mask(in[8] != 0.f), Flt(in[9]), Flt(in[10]),
                                                   in[11] != 0.f, in[12] != 0.f);
            out[0] = r[0][0]; out[1] = r[1][0]; out[2] = r[2][0]; return 3; }
        case KAT_DIST_PARAMS: { auto [ax, ay] = calc_dist_params<Flt>(Flt(in[0]), Flt(in[1]), in[2] != 0.f); out[0] = f1(ax); out[1] = f1(ay); return 2; }
        case KAT_COORD_SYSTEM: { auto [s, t] = coordinate_system(v3(in)); put3(out, s); put3(out + 3, t); return 6; }
        case KAT_ERFINV: out[0] = f1(dr::erfinv(Flt(in[0]))); return 1;
    }
    return -1;
}
