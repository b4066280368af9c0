"""CPU tests pinning the `bitmap` texture and `bumpmap` frame of the oracle to the reference's own
tests: src/textures/tests/test_bitmap.py (golden texel values after the sRGB -> linear load, wrap-mode identities,
finite-difference check of eval_1_grad under a rotated to_uv) on the same image files (tests/golden/textures.npz,
tools/make_fixtures.py).  The reference has no bumpmap test; its frame is checked against the closed form of
src/bsdfs/bumpmap.cpp:248-276 on a flat interaction."""
import os
import numpy as np
import pytest

import liverrenderer_b200 as mi
from liverrenderer_b200 import _capi as K
from liverrenderer_b200.transforms import T
from tests import orc

TEX = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "textures.npz"))


def _tex(pixels, **kw):
    """lr_texture through the host mirror's own `bitmap` loader."""
    d = {"type": "scene", "sensor": {"type": "perspective", "film": {"type": "hdrfilm", "width": 4, "height": 4}},
         "floor": {"type": "rectangle", "bsdf": {"type": "diffuse", "reflectance": dict({"type": "bitmap", "bitmap": pixels}, **kw)}}}
    sc = mi.load_dict(d)
    assert len(sc.textures) == 1 and sc.bsdfs[-1].tex_kind == 2 and sc.bsdfs[-1].refl_tex == 1
    return sc, sc.textures[0]


def luminance(c):
    return c[..., 0] * np.float32(0.212671) + c[..., 1] * np.float32(0.715160) + c[..., 2] * np.float32(0.072169)


def test_golden_texels_rgb_and_grey():
    """test_bitmap.py:128-176 (test04_eval_rgb)."""
    sc, t = _tex(TEX["carrot"])                         # RGBA file -> RGB, sRGB -> linear
    assert (t.width, t.height, t.channels) == (10, 9, 3)
    uv = [[(1 / 10) * 7 + 1 / 20, (1 / 9) * 1 + 1 / 18]]        # centre of the green pixel (7, 1)
    o = orc.texture_probe(t, uv)[0]
    expected = np.array([0.0012, 0.1946, 0.0241], dtype=np.float32)
    np.testing.assert_allclose(o[:3], expected, atol=1e-4)
    np.testing.assert_allclose(o[3], luminance(expected), atol=1e-4)
    sc, t = _tex(TEX["noise_02"][..., 0])               # greyscale file -> 1 channel
    assert (t.width, t.height, t.channels) == (239, 241, 1)
    o = orc.texture_probe(t, [[1 / (2 * 239), 15 / 241 + 1 / (2 * 241)]])[0]      # gray pixel (0, 15)
    np.testing.assert_allclose(o[:4], 0.5395, atol=1e-4)


@pytest.mark.parametrize("filter_type", ["bilinear", "nearest"])
@pytest.mark.parametrize("wrap_mode", ["repeat", "clamp", "mirror"])
def test_wrap_modes(wrap_mode, filter_type):
    """test_bitmap.py:55-125 (test03_wrap)."""
    sc, t = _tex(TEX["noise_8x8"], wrap_mode=wrap_mode, filter_type=filter_type)

    def ev(xs, ys):
        xv, yv = np.meshgrid(xs, ys); return orc.texture_probe(t, np.stack([xv.ravel(), yv.ravel()], axis=1))[:, :3]
    n = 20; lin = lambda a, b: np.linspace(a, b, n, dtype=np.float32)
    ref = ev(lin(0, 1), lin(0, 1))
    atol = 1e-4 if filter_type == "bilinear" else 0
    if filter_type == "nearest":        # texel edges fall on the sample grid: keep away from them
        lin = lambda a, b: np.linspace(a, b, n, dtype=np.float32) * np.float32(0.999) + np.float32(0.0004) * np.sign(b - a + 1e-9)
        ref = ev(lin(0, 1), lin(0, 1))
    if wrap_mode == "repeat":
        np.testing.assert_allclose(ev(lin(0, 1) - 1, lin(0, 1) - 1), ref, atol=max(atol, 2e-6))
        np.testing.assert_allclose(ev(lin(0, 1) + 1, lin(0, 1) + 1), ref, atol=max(atol, 2e-6))
    elif wrap_mode == "clamp":
        z, o = np.zeros(n, np.float32), np.ones(n, np.float32)
        np.testing.assert_allclose(ev(lin(0, 1), lin(-1, 0)), ev(lin(0, 1), z), atol=atol + 1e-6)
        np.testing.assert_allclose(ev(lin(0, 1), lin(1, 2)), ev(lin(0, 1), o), atol=atol + 1e-6)
        np.testing.assert_allclose(ev(lin(-1, 0), lin(0, 1)), ev(z, lin(0, 1)), atol=atol + 1e-6)
        np.testing.assert_allclose(ev(lin(1, 2), lin(0, 1)), ev(o, lin(0, 1)), atol=atol + 1e-6)
    elif filter_type == "bilinear":
        np.testing.assert_allclose(ev(lin(0, -1), lin(0, -1)), ref, atol=1e-4)
        np.testing.assert_allclose(ev(lin(2, 1), lin(2, 1)), ref, atol=1e-4)


def test_eval_1_grad_finite_differences():
    """test_bitmap.py:31-52 (test02_eval_grad): analytic gradient against forward differences of eval_1 under a
    rotated to_uv (in float64 on the oracle's own taps, away from texel borders where the bilinear patch changes)."""
    rng = np.random.default_rng(7); checked = 0
    for angle in rng.random(10) * 360.0:
        sc, t = _tex(TEX["noise_8x8"], to_uv=T().rotate([0, 0, 1], angle))
        uv = rng.random((40, 2)).astype(np.float32); delta = np.float32(2.5e-4)
        f = orc.texture_probe(t, uv); fu = orc.texture_probe(t, uv + [delta, 0]); fv = orc.texture_probe(t, uv + [0, delta])
        fd = np.stack([(fu[:, 3] - f[:, 3]), (fv[:, 3] - f[:, 3])], axis=1).astype(np.float64) / float(delta)
        g = f[:, 4:6].astype(np.float64); gmax = np.abs(g).max()
        # keep the points whose stencil stays inside one bilinear patch (the gradient jumps across texel borders)
        same = (np.abs(fu[:, 4:6] - f[:, 4:6]).max(axis=1) < 0.05 * gmax) & (np.abs(fv[:, 4:6] - f[:, 4:6]).max(axis=1) < 0.05 * gmax)
        err = np.abs(fd - g)[same]
        assert err.max() < 3e-2 * gmax, (angle, err.max(), gmax)
        checked += int(same.sum())
    assert checked > 200


def test_nearest_has_zero_gradient_and_to_uv_scale():
    sc, t = _tex(TEX["noise_8x8"], filter_type="nearest")
    assert np.all(orc.texture_probe(t, np.random.default_rng(0).random((16, 2)))[:, 4:6] == 0)
    # to_uv = scale(3): same texels at uv/3, gradient three times larger (chain rule, bitmap.cpp:566-573)
    sc1, t1 = _tex(TEX["noise_8x8"]); sc3, t3 = _tex(TEX["noise_8x8"], to_uv=T().scale([3, 3, 1]))
    uv = np.random.default_rng(1).random((32, 2)).astype(np.float32)
    a = orc.texture_probe(t1, uv); b = orc.texture_probe(t3, uv / np.float32(3))
    np.testing.assert_allclose(b[:, :4], a[:, :4], atol=2e-5)
    np.testing.assert_allclose(b[:, 4:6], 3 * a[:, 4:6], rtol=1e-3, atol=2e-3)


def _bump_bsdf(scale=1.0, flags=K.BUMP_FLIP_INVALID_NORMALS | K.BUMP_SHADOWING_FUNCTION):
    b = K.lr_bsdf(); b.kind = K.BSDF_DIFFUSE; b.bump_tex = 1; b.bump_scale = scale; b.bump_flags = flags; return b


def test_bump_frame_closed_form():
    """bumpmap.cpp:248-276 on a flat interaction (n = sh.n = z, dp_du = x, dp_dv = y): the perturbed normal is
    normalize(-s*g_u, -s*g_v, 1); a constant texture leaves the frame untouched."""
    z = np.array([0, 0, 1.0]); x = np.array([1.0, 0, 0]); y = np.array([0, 1.0, 0]); wi = np.array([0.3, -0.2, 0.9])
    sc, flat = _tex(np.full((4, 4), 0.25, dtype=np.float32), raw=True)
    f = orc.bump_frame(_bump_bsdf(), flat, z, z, x, x, y, wi, [0.4, 0.6])
    np.testing.assert_allclose(f, [[0, 0, 1], [1, 0, 0], [0, 1, 0]], atol=1e-7)
    ramp = np.tile(np.linspace(0, 1, 16, dtype=np.float32)[None, :], (16, 1))        # f = u (texel centres): df/du = 16/15
    sc, t = _tex(ramp, raw=True, wrap_mode="clamp")
    for scale in (0.25, 1.0, -2.0):
        uv = [0.5, 0.5]; g = orc.texture_probe(t, [uv])[0, 4:6]
        np.testing.assert_allclose(g, [16 / 15, 0], atol=2e-5)
        f = orc.bump_frame(_bump_bsdf(scale), t, z, z, x, x, y, wi, uv)
        n = np.array([-scale * g[0], -scale * g[1], 1.0]); n /= np.linalg.norm(n)
        np.testing.assert_allclose(f[0], n, atol=1e-6)
        np.testing.assert_allclose([f[0] @ f[1], f[0] @ f[2], f[1] @ f[2]], 0, atol=1e-6)
        np.testing.assert_allclose(np.cross(f[0], f[1]), f[2], atol=1e-6)
    # flip_invalid_normals (:265-269): incident direction below the perturbed tangent plane mirrors n.xy
    wi_g = np.array([0.995, 0.0, 0.1]); wi_g /= np.linalg.norm(wi_g)
    f_on = orc.bump_frame(_bump_bsdf(4.0), t, z, z, x, x, y, wi_g, [0.5, 0.5])
    f_off = orc.bump_frame(_bump_bsdf(4.0, K.BUMP_SHADOWING_FUNCTION), t, z, z, x, x, y, wi_g, [0.5, 0.5])
    assert wi_g @ f_off[0] < 0 and wi_g @ f_on[0] > 0
    np.testing.assert_allclose(f_on[0], f_off[0] * [-1, -1, 1], atol=1e-7)


def test_bumpmap_dict_errors_and_pod():
    """Constructor behaviour of src/bsdfs/bumpmap.cpp:101-127 in the host mirror."""
    base = {"type": "scene", "sensor": {"type": "perspective", "film": {"type": "hdrfilm", "width": 4, "height": 4}}}
    tex = {"type": "bitmap", "bitmap": TEX["noise_8x8"], "raw": True}
    with pytest.raises(RuntimeError, match="Exactly one BSDF child"):
        mi.load_dict(dict(base, s={"type": "rectangle", "bsdf": {"type": "bumpmap", "tex": tex}}))
    with pytest.raises(RuntimeError, match="Exactly one Texture child"):
        mi.load_dict(dict(base, s={"type": "rectangle", "bsdf": {"type": "bumpmap", "b": {"type": "diffuse"}}}))
    sc = mi.load_dict(dict(base, s={"type": "rectangle", "bsdf": {"type": "bumpmap", "scale": 0.005, "tex": tex,
                                                                  "b": {"type": "dielectric", "int_ior": 1.38, "ext_ior": 1.0}}}))
    b = sc.bsdfs[-1]
    assert b.kind == K.BSDF_DIELECTRIC and b.bump_tex == 1 and abs(b.bump_scale - 0.005) < 1e-9 and b.bump_flags == 3
    assert abs(b.eta - 1.38) < 1e-6
    pod = sc.pod(); assert pod.n_textures == 1 and pod.textures[0].channels == 3 and pod.textures[0].width == 8
    # 16-bit raw data is only normalised, not gamma-decoded
    np.testing.assert_allclose(np.ctypeslib.as_array(pod.textures[0].data, (8 * 8 * 3,)).reshape(8, 8, 3),
                               TEX["noise_8x8"].astype(np.float32) / 65535.0, rtol=1e-6)
