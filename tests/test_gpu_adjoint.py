"""prbvolpath on the GPU: primal parity, adjoint vs the oracle adjoint, and the finite-difference
check of the density-grid gradient (the reference itself does not pin prbvolpath gradients,
SURVEY.md 8c: FD with common random numbers is the gate the north star names, 1e-3 relative
on the loss derivative)."""
import numpy as np
import pytest

import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes
from tests import orc, golden_scenes as G

pytestmark = pytest.mark.gpu


def _small_inverse(res=24, spp=16, grid_res=8, seed_density=0):
    rng = np.random.default_rng(12345 + seed_density)
    dens = (0.02 + 0.05 * rng.random((grid_res, grid_res, grid_res, 1))).astype(np.float32)
    d = scenes.sphere_liver_inverse(res=res, spp=spp, grid_res=grid_res, density=dens)
    # fixed majorant (`max_value`, src/volumes/grid.cpp:303-306): perturbing the grid then leaves the
    # free-flight samples untouched, so finite differences see common random numbers
    d["sphere"]["interior"]["sigma_t"]["max_value"] = 0.08
    return d


def test_prb_primal_per_path_matches_oracle():
    sc = mi.load_dict(_small_inverse(res=32, spp=8)); n = 32 * 32 * 8
    gpu = mi.Context(0).upload(sc).sample_lanes(0, 8, 0, n); cpu = orc.OracleScene(sc).sample_lanes(0, 8, 0, n)
    err = np.abs(gpu[:, :3] - cpu[:, :3]) / np.maximum(np.abs(cpu[:, :3]), 1e-3)
    assert (err.max(axis=1) < 1e-3).mean() > 0.97
    assert abs(gpu[:, :3].mean() - cpu[:, :3].mean()) / cpu[:, :3].mean() < 1e-2


@pytest.mark.parametrize("name", ["test_hetero_sigmat", "test_homo_null_envmap"])
def test_prb_primal_passes_reference_ztest(name):
    """INTEGRATOR_MAPPING (src/render/tests/test_renders.py:45-49): every *_volpath scene is also
    rendered with prbvolpath against the same golden EXRs."""
    spp = int(2e6) // (48 * 48)
    sc = mi.load_dict(G.scene(name, integrator="prbvolpath", spp=spp))
    img = mi.Context(0).upload(sc).render(seed=0, spp=spp)
    frac, ok = G.reference_accepts(img, name, spp)
    assert ok, f"{name} (prbvolpath primal): only {frac:.4f} accepted"


def test_adjoint_matches_oracle_adjoint():
    sc = mi.load_dict(_small_inverse(res=24, spp=16))
    rng = np.random.default_rng(7); g = rng.standard_normal((24, 24, 3)).astype(np.float32)
    ctx = mi.Context(0).upload(sc); ctx.grid_grad_zero(0); ctx.render_backward(g, seed=11, spp=16)
    o = orc.OracleScene(sc); o.grid_grad_zero(0); o.render_backward(g, seed=11, spp=16)
    a, b = ctx.grid_grad_read(0), o.grid_grad_read(0)
    assert np.isfinite(a).all()
    assert np.linalg.norm(a - b) / np.linalg.norm(b) < 2e-2
    # accumulation: a second call adds
    ctx.render_backward(g, seed=11, spp=16)
    np.testing.assert_allclose(ctx.grid_grad_read(0), 2 * a, rtol=1e-3, atol=1e-6 * np.abs(a).max())


def test_adjoint_lane_split_sums_to_whole():
    """Two ranks each replay half of the lanes; the all-reduce (sum) of their gradient buffers must
    equal the single-GPU gradient."""
    sc = mi.load_dict(_small_inverse(res=16, spp=8)); n = 16 * 16 * 8
    g = np.ones((16, 16, 3), dtype=np.float32)
    ctx = mi.Context(0).upload(sc)
    ctx.grid_grad_zero(0); ctx.render_backward(g, seed=3, spp=8); whole = ctx.grid_grad_read(0)
    ctx.grid_grad_zero(0); ctx.render_backward(g, seed=3, spp=8, begin=0, end=n // 2); h0 = ctx.grid_grad_read(0)
    ctx.grid_grad_zero(0); ctx.render_backward(g, seed=3, spp=8, begin=n // 2, end=n); h1 = ctx.grid_grad_read(0)
    np.testing.assert_allclose(h0 + h1, whole, rtol=2e-4, atol=1e-6 * np.abs(whole).max())


def test_gradient_finite_difference():
    """Directional derivative of loss(render(sigma_t + eps * dir)) by central differences with common
    random numbers (same seed, fixed majorant) vs <grad, dir> from render_backward.  Delta tracking makes
    the same-seed render piecewise constant in sigma_t per path (null/real decisions flip), so the FD of
    a finite-sample render carries O(sqrt(1/(eps*N))) noise: N = 16.8M paths here."""
    res, spp, gr = 16, 65536, 4
    d = _small_inverse(res=res, spp=spp, grid_res=gr)
    dens0 = np.array(d["sphere"]["interior"]["sigma_t"]["grid"], dtype=np.float32)
    sc = mi.load_dict(d); ctx = mi.Context(0).upload(sc)
    rng = np.random.default_rng(12345)
    target = rng.random((res, res, 3)).astype(np.float32) * 0.5
    seed = 9

    def loss_and_img(dens):
        ctx.grid_update(0, np.ascontiguousarray(dens, dtype=np.float32))
        img = ctx.render(seed=seed, spp=spp)
        return float(((img.astype(np.float64) - target) ** 2).mean()), img

    l0, img0 = loss_and_img(dens0)
    grad_in = (2.0 * (img0.astype(np.float64) - target) / img0.size).astype(np.float32)
    ctx.grid_grad_zero(0)
    ctx.render_backward(grad_in, seed=seed, spp=spp)           # same seed => the adjoint replays the primal paths
    grad = ctx.grid_grad_read(0).reshape(dens0.shape).astype(np.float64)
    assert np.isfinite(grad).all() and np.abs(grad).max() > 0

    dirs = [np.ones_like(dens0, dtype=np.float64)]
    zz = np.arange(gr)[:, None, None, None] / (gr - 1.0)
    dirs.append(np.broadcast_to(zz, dens0.shape).astype(np.float64))                 # linear ramp along z
    dirs.append(np.broadcast_to(1.0 - zz.transpose(1, 0, 2, 3), dens0.shape).astype(np.float64))   # ramp along y
    report = []
    for k, m in enumerate(dirs):
        eps = 1e-3
        lp, _ = loss_and_img(dens0 + eps * m); lm, _ = loss_and_img(dens0 - eps * m)
        fd = (lp - lm) / (2 * eps); an = float((grad * m).sum())
        report.append((k, fd, an, abs(fd - an) / max(abs(fd), abs(an))))
    ctx.grid_update(0, dens0)
    print("FD check (dir, fd, analytic, rel):", report)
    for k, fd, an, rel in report:
        assert rel < 2e-2, report


def test_traverse_update_recomputes_majorant():
    d = _small_inverse(res=16, spp=4); d["sphere"]["interior"]["sigma_t"].pop("max_value")
    sc = mi.load_dict(d)
    params = mi.traverse(sc)
    key = [k for k in params.keys() if k.endswith("sigma_t.data")][0]
    img0 = mi.render(sc, spp=4, seed=1)
    params[key] = params[key] * 3.0; params.update()
    img1 = mi.render(sc, spp=4, seed=1)
    assert np.isfinite(img1).all() and not np.allclose(img0, img1)
    o = orc.OracleScene(sc)           # oracle built from the updated host array
    ref = o.render(seed=1, spp=4)
    assert np.median(np.abs(img1 - ref) / np.maximum(np.abs(ref), 1e-2)) < 1e-4
