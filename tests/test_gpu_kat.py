"""Device known-answer tests: the reference's function-level tests restated on the CUDA code itself through the
lr_kat_* probes of the C ABI (one query per thread), next to the same probes of the oracle.  What tests/test_oracle_kat.py
pins on the oracle is pinned here on the product: principled 20 + 20 goldens, diffuse closed form, sample/eval/pdf
consistency for every BSDF configuration (GGX and Beckmann, visible and non-visible normal sampling, every principled
lobe), phase functions, gridvolume trilinear / nearest x clamp / repeat / mirror, the envmap warp."""
import ctypes as C
import math

import numpy as np
import pytest

import liverrenderer_b200 as mi
from liverrenderer_b200 import _capi as K, scenes
from tests import orc
from tests.test_oracle_kat import PRINCIPLED_PDF, PRINCIPLED_EVAL, _bsdf_pod, _grid_scene, _sphere_quadrature, _p

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    return mi.Context(0)


def _orc_eval_pdf(b, wi, wo):
    L = orc.lib(); ev = np.zeros((len(wi), 3), np.float32); pdf = np.zeros(len(wi), np.float32)
    for i in range(len(wi)):
        e = np.zeros(3, np.float32); p = np.zeros(1, np.float32)
        L.orc_kat_bsdf_eval_pdf(C.byref(b), _p(np.ascontiguousarray(wi[i], np.float32)), _p(np.ascontiguousarray(wo[i], np.float32)), _p(e), _p(p))
        ev[i] = e; pdf[i] = p[0]
    return ev, pdf


def test_principled_goldens_on_device(ctx):
    """src/bsdfs/tests/test_principled.py:124-205 on CUDA: 20 pdf + 20 eval values of the 10-parameter principled BSDF,
    wi = (1, 0, 1) unnormalised as in the reference test; tolerance = dr.allclose defaults (rtol 1e-5, atol 1e-8)."""
    b = _bsdf_pod({"type": "principled", "metallic": 0.3, "spec_tint": 0.6, "specular": 0.4, "anisotropic": 0.3, "clearcoat": 0.3,
                   "clearcoat_gloss": 0.5, "sheen": 0.9, "sheen_tint": 0.9, "spec_trans": 0.5, "flatness": 0.5})
    theta = np.arange(20) / 19.0 * (math.pi / 2)
    wo = np.stack([np.sin(theta), np.zeros(20), np.cos(theta)], -1); wi = np.tile([1.0, 0.0, 1.0], (20, 1))
    ev, pdf = ctx.kat_bsdf_eval_pdf(b, wi, wo)
    assert np.allclose(pdf, PRINCIPLED_PDF, rtol=1e-5, atol=1e-8), pdf
    assert np.allclose(ev[:, 0], PRINCIPLED_EVAL, rtol=1e-5, atol=1e-8), ev[:, 0]


def test_diffuse_closed_form_on_device(ctx):
    """src/bsdfs/tests/test_diffuse.py:14-37"""
    b = _bsdf_pod({"type": "diffuse"})
    theta = np.arange(20) / 19.0 * (math.pi / 2); wo = np.stack([np.sin(theta), np.zeros(20), np.cos(theta)], -1)
    ev, pdf = ctx.kat_bsdf_eval_pdf(b, np.tile([0.0, 0.0, 1.0], (20, 1)), wo)
    assert np.allclose(pdf, np.cos(theta) / math.pi, atol=1e-6)
    assert np.allclose(ev, (0.5 * np.cos(theta) / math.pi)[:, None], atol=1e-6)


BSDFS = {
    "rd_beckmann": {"type": "roughdielectric", "alpha": 0.3, "distribution": "beckmann", "int_ior": 1.5, "ext_ior": 1.0},
    "rd_ggx": {"type": "roughdielectric", "alpha": 0.2, "distribution": "ggx", "int_ior": 1.38, "ext_ior": 1.0},
    "rd_beckmann_nonvisible": {"type": "roughdielectric", "alpha": 0.25, "distribution": "beckmann", "int_ior": 1.5, "ext_ior": 1.0, "sample_visible": False},
    "rd_ggx_nonvisible_aniso": {"type": "roughdielectric", "alpha_u": 0.15, "alpha_v": 0.35, "distribution": "ggx", "int_ior": 1.33, "ext_ior": 1.0, "sample_visible": False},
    "principled_trans": {"type": "principled", "base_color": {"type": "rgb", "value": [0.8, 0.3, 0.25]}, "roughness": 0.3, "spec_trans": 0.9, "eta": 1.38},
    "principled_clearcoat_sheen": {"type": "principled", "base_color": {"type": "rgb", "value": [0.6, 0.5, 0.2]}, "roughness": 0.4, "clearcoat": 0.8,
                                   "clearcoat_gloss": 0.7, "sheen": 0.6, "sheen_tint": 0.4, "specular": 0.6},
    "principled_metal_aniso": {"type": "principled", "base_color": {"type": "rgb", "value": [0.9, 0.7, 0.3]}, "roughness": 0.35, "metallic": 0.8,
                               "anisotropic": 0.6, "spec_tint": 0.5, "specular": 0.5},
    "principled_flat": {"type": "principled", "base_color": {"type": "rgb", "value": [0.5, 0.6, 0.7]}, "roughness": 0.5, "flatness": 0.7, "spec_trans": 0.3,
                        "specular": 0.5},
    "principled_all": {"type": "principled", "metallic": 0.3, "spec_tint": 0.6, "specular": 0.4, "anisotropic": 0.3, "clearcoat": 0.3,
                       "clearcoat_gloss": 0.5, "sheen": 0.9, "sheen_tint": 0.9, "spec_trans": 0.5, "flatness": 0.5},
    "diffuse": {"type": "diffuse"},
    "dielectric": {"type": "dielectric", "int_ior": 1.38, "ext_ior": 1.0},
}


@pytest.mark.parametrize("name", list(BSDFS))
def test_bsdf_device_equals_oracle_and_is_self_consistent(ctx, name):
    """(1) CUDA sample / eval / pdf equal the oracle's on the same inputs (the oracle is pinned to the reference's goldens);
    (2) src/bsdfs/tests/test_rough_dielectric.py:209-229 on the device: sample() weight == eval()/pdf(), sample pdf == pdf().
    Both sides of the surface (wi.z > 0 and < 0) are probed."""
    b = _bsdf_pod(BSDFS[name]); rng = np.random.default_rng(1); n = 2000
    wi = rng.standard_normal((n, 3)); wi[:, 2] = np.sign(wi[:, 2]) * (np.abs(wi[:, 2]) + 0.1); wi /= np.linalg.norm(wi, axis=1, keepdims=True)
    u = rng.random((n, 3)).astype(np.float32); wi = wi.astype(np.float32)
    s = ctx.kat_bsdf_sample(b, wi, u)
    assert np.isfinite(s["wo"]).all() and np.isfinite(s["weight"]).all()
    # oracle on the same inputs
    L = orc.lib(); o_out = np.zeros((n, 9), np.float32); o_t = np.zeros(n, np.uint32)
    for i in range(n):
        o_t[i] = L.orc_kat_bsdf_sample(C.byref(b), _p(np.ascontiguousarray(wi[i])), float(u[i, 0]), float(u[i, 1]), float(u[i, 2]), _p(o_out[i]))
    same_lobe = o_t == s["sampled_type"]
    assert same_lobe.mean() > 0.995, same_lobe.mean()          # a lobe pick can flip where u sits within an ulp of the Fresnel term
    m = same_lobe & (o_out[:, 3] > 1e-4)
    assert np.allclose(s["wo"][m], o_out[m, 0:3], atol=2e-4)
    assert np.allclose(s["pdf"][m], o_out[m, 3], rtol=2e-3, atol=1e-5)
    assert np.allclose(s["weight"][m], o_out[m, 5:8], rtol=2e-3, atol=1e-5)
    if name == "dielectric":
        return                                                  # delta lobes: eval / pdf are 0 by definition
    ok = m & (s["pdf"] > 1e-4)
    t = s["sampled_type"]; prod = s["wo"][:, 2] * wi[:, 2]
    ok &= ~(((t & 0x6) != 0) & (prod <= 0)) & ~(((t & 0x8) != 0) & (prod >= 0))        # roughdielectric.cpp:377-380
    ev, pdf = ctx.kat_bsdf_eval_pdf(b, wi[ok], s["wo"][ok])
    o_ev, o_pdf = _orc_eval_pdf(b, wi[ok][:200], s["wo"][ok][:200])
    assert np.allclose(ev[:200], o_ev, rtol=2e-3, atol=1e-6) and np.allclose(pdf[:200], o_pdf, rtol=2e-3, atol=1e-6)
    dp = np.abs(pdf - s["pdf"][ok]) / np.maximum(s["pdf"][ok], 1.0)          # pdf() rebuilds the half vector from (wi, wo): fp32 round trip, worst for small alpha
    assert np.quantile(dp, 0.98) <= 2e-3 and (dp.max() <= 5e-2 or "nonvisible" in name), (np.quantile(dp, 0.98), dp.max())
    if "nonvisible" not in name:
        # (with sample_visible = false the reference samples a roughened copy of the distribution, alpha * (1.2 - 0.2 sqrt|cos|), but weights
        # with the plain one, roughdielectric.cpp:268-276,338-343: weight != eval / pdf there by construction -- device == oracle is the check)
        assert np.allclose(ev / pdf[:, None], s["weight"][ok], rtol=5e-3, atol=1e-4)
    assert ok.sum() > n // 3


@pytest.mark.parametrize("kind,g", [(K.PHASE_ISOTROPIC, 0.0), (K.PHASE_HG, 0.6), (K.PHASE_HG, -0.6), (K.PHASE_HG, 0.8)])
def test_phase_functions_on_device(ctx, kind, g):
    """src/phase/tests/test_hg.py:11-24 / test_isotropic.py:11-24 on CUDA: the pdf integrates to 1, sampled pdf == eval,
    the mean cosine of the sampled directions is g, and values equal the oracle's."""
    rng = np.random.default_rng(5); n = 20000
    wi = np.tile(np.array([0, 0, 1], np.float32), (n, 1)); u = rng.random((n, 2)).astype(np.float32)
    wo, pdf, ev = ctx.kat_phase(kind, g, wi, u)
    assert np.allclose(np.linalg.norm(wo, axis=1), 1.0, atol=1e-4)
    assert np.allclose(pdf, ev, rtol=1e-4, atol=1e-7)
    assert abs(float(np.mean(-wo[:, 2])) - g) < 0.01
    assert abs(float(np.mean(1.0 / (4 * math.pi * pdf.astype(np.float64)))) - 1.0) < 0.05      # E_p[1 / (4 pi p)] = 1: the pdf is normalised
    L = orc.lib(); out = np.zeros(4, np.float32)
    for i in range(0, 200):
        L.orc_kat_phase_sample(kind, g, _p(wi[i]), float(u[i, 0]), float(u[i, 1]), _p(out))
        assert np.allclose(out[:3], wo[i], atol=2e-4) and abs(out[3] - pdf[i]) <= 1e-3 * max(1.0, out[3])      # glibc vs CUDA sincosf


@pytest.mark.parametrize("wrap", ["clamp", "repeat", "mirror"])
@pytest.mark.parametrize("filt", ["trilinear", "nearest"])
def test_gridvolume_lookup_on_device(filt, wrap):
    """src/volumes/tests/test_grid.py:43-118 on CUDA for every filter x wrap mode: equal to the oracle (itself pinned to the
    reference's goldens) at random points inside and OUTSIDE the unit cube, the reference's golden lookups, and the fat-voxel
    path (trilinear + clamp) bit-identical to the reference-layout path."""
    rng = np.random.default_rng(3); g = rng.random((5, 6, 7)).astype(np.float32)
    d = mi.cornell_box(); d["integrator"] = {"type": "volpath", "max_depth": 4}
    d["small-box"]["interior"] = {"type": "heterogeneous", "albedo": {"type": "rgb", "value": [0.5, 0.5, 0.5]},
                                  "sigma_t": {"type": "gridvolume", "grid": g, "wrap_mode": wrap, "filter_type": filt}}
    sc = mi.load_dict(d); ctx = mi.Context(0).upload(sc); o = orc.OracleScene(sc)
    p = np.concatenate([rng.random((400, 3)), rng.uniform(-2.2, 3.1, size=(400, 3)), [[0, 0, 0], [1, 1, 1], [0.5, 0.5, 0.0], [1 / 3, 1 / 3, 1 / 3]]]).astype(np.float32)
    dev = ctx.kat_grid_eval(0, p, use_fat=True)[:, 0]; plain = ctx.kat_grid_eval(0, p, use_fat=False)[:, 0]
    ref = np.zeros(len(p), np.float32); out = np.zeros(3, np.float32)
    for i in range(len(p)):
        o.L.orc_kat_grid_eval(o.h, 0, _p(np.ascontiguousarray(p[i])), _p(out)); ref[i] = out[0]
    assert np.array_equal(dev, plain)                                          # fat-voxel path == reference-layout path, bit for bit
    assert np.allclose(dev, ref, atol=2e-6), np.abs(dev - ref).max()
    if wrap == "clamp" and filt == "trilinear":                                # the reference's own golden lookups
        gg = np.ones((3, 3, 3), dtype=np.float32); gg[0, 0, 0] = 0.0; gg[1, 1, 1] = 0.5
        c2 = mi.Context(0).upload(_grid_scene(gg))
        v = c2.kat_grid_eval(0, [[0, 0, 0], [1 / 3, 1 / 3, 1 / 3]])[:, 0]
        assert np.allclose(v, [0.0, (6 * 1 + 1 * 0.5 + 1 * 0.0) / 8])
        gb = np.ones((3, 8, 5), dtype=np.float32); gb[0, 0, 0] = 0.0; gb[1, 1, 1] = 0.5; gb[:, 7, :] = 0.55; gb[0, :, :] = 0.32
        c3 = mi.Context(0).upload(_grid_scene(gb))
        assert np.allclose(c3.kat_grid_eval(0, [[0.2, 1.0, 0.7], [0.5, 0.5, 0.0]])[:, 0], [0.55, 0.32])


def test_envmap_warp_on_device():
    """src/emitters/tests/test_envmap.py on CUDA with the C3 environment map: the hierarchical sample warp equals the oracle's,
    pdf_direction(sampled d) == sampled pdf, weight == eval / pdf, and the pdf integrates to 1 over the sphere."""
    sc = mi.load_dict(scenes.liver_single_mesh(res=16, spp=1, grid_res=8)); ctx = mi.Context(0).upload(sc); o = orc.OracleScene(sc)
    rng = np.random.default_rng(7); n = 4000
    ref_p = rng.uniform(-60, -20, size=(n, 3)).astype(np.float32); u = rng.random((n, 2)).astype(np.float32)
    d, pdf, w = ctx.kat_emitter_sample(ref_p, u)
    assert np.allclose(np.linalg.norm(d, axis=1), 1.0, atol=1e-4) and (pdf > 0).mean() > 0.99
    out = np.zeros(7, np.float32)
    for i in range(300):
        o.L.orc_kat_env_sample(o.h, _p(np.ascontiguousarray(ref_p[i])), float(u[i, 0]), float(u[i, 1]), _p(out))
        assert np.allclose(out[:3], d[i], atol=2e-5) and abs(out[3] - pdf[i]) <= 1e-4 * max(out[3], 1e-3) and np.allclose(out[4:7], w[i], rtol=1e-3, atol=1e-5)
    e, pdf2 = ctx.kat_env_eval_pdf(d)
    m = pdf > 1e-6
    assert np.allclose(pdf2[m], pdf[m], rtol=2e-3)
    assert np.allclose((e[m] / pdf2[m, None]), w[m], rtol=5e-3, atol=1e-4)
    q, dw = _sphere_quadrature(512, 1024)
    _, pq = ctx.kat_env_eval_pdf(q)
    assert abs(float(pq.astype(np.float64).sum() * dw) - 1.0) < 2e-2
