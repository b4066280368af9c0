/* kat_fn.h -- function ids of the generic known-answer probe `*_kat_fn(fn, in, out)` that BOTH the oracle (orc_kat_fn,
   oracle.cpp: the hand restatement) and oracle/_ref/libref_kat.so (ref_kat_fn, ref_kat/ref_kat.cpp: the REFERENCE'S OWN
   header-only code compiled for Float = float) export, so tests/test_oracle_ref.py can compare them bit for bit.
   TEST INFRASTRUCTURE ONLY.  `in`: up to 24 floats (integers passed as bit patterns), `out`: up to 8 floats.            */
#pragma once
enum {
    KAT_TEA32 = 0,            /* in: v0, v1, rounds (bits)                    out: 2 x uint32 bits                        */
    KAT_TEA_FLOAT32,          /* in: v0, v1, rounds (bits)                    out: 1                                      */
    KAT_PCG32,                /* in: initstate lo, hi, initseq lo, hi (bits)  out: 4 x next_float32, 4 x next_uint32 bits */
    KAT_WARP_DISK,            /* in: u(2)                                     out: 2                                      */
    KAT_WARP_COS_HEMI,        /* in: u(2)                                     out: v(3), pdf(v)                           */
    KAT_WARP_SPHERE,          /* in: u(2)                                     out: 3                                      */
    KAT_WARP_TRIANGLE,        /* in: u(2)                                     out: 2                                      */
    KAT_FRESNEL,              /* in: cos_theta_i, eta                         out: F, cos_theta_t, eta_it, eta_ti         */
    KAT_REFLECT,              /* in: wi(3), m(3)                              out: 3                                      */
    KAT_REFRACT,              /* in: wi(3), m(3), cos_theta_t, eta_ti         out: 3                                      */
    KAT_MF_EVAL,              /* in: type, au, av, sample_visible, m(3)       out: 1        (type, sample_visible as floats) */
    KAT_MF_PDF,               /* in: type, au, av, sv, wi(3), m(3)            out: 1                                      */
    KAT_MF_SAMPLE,            /* in: type, au, av, sv, wi(3), u(2)            out: m(3), pdf                              */
    KAT_MF_G1,                /* in: type, au, av, sv, v(3), m(3)             out: 1                                      */
    KAT_MF_G,                 /* in: type, au, av, sv, wi(3), wo(3), m(3)     out: 1                                      */
    KAT_GTR1,                 /* in: alpha, m(3), u(2)                        out: eval(m), pdf(m), sample(u)(3)          */
    KAT_SMITH_GGX1,           /* in: v(3), wh(3), alpha, wo(3)                out: smith_ggx1(v, wh), clearcoat_G(v, wo, wh) */
    KAT_SCHLICK,              /* in: cos, R0, eta                             out: schlick_weight, calc_schlick, schlick_R0_eta */
    KAT_MAC_MIC,              /* in: m(3), wi(3), wo(3), cos_theta_i          out: reflection case, refraction case (0/1) */
    KAT_PRINCIPLED_FRESNEL,   /* in: F_diel, metallic, spec_tint, base(3), lum, cos_theta_i, front, bsdf, eta, has_metallic, has_spec_tint   out: 3 */
    KAT_DIST_PARAMS,          /* in: anisotropic, roughness, has_aniso        out: ax, ay                                 */
    KAT_COORD_SYSTEM,         /* in: n(3)                                     out: s(3), t(3)                             */
    KAT_ERFINV,               /* in: x                                        out: 1                                      */
    KAT_COUNT
};
