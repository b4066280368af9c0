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
    print("adjoint vs oracle adjoint: rel L2", np.linalg.norm(a - b) / np.linalg.norm(b), "sum ratio", a.sum() / b.sum())
    assert np.linalg.norm(a - b) / np.linalg.norm(b) < 2e-2 and abs(a.sum() / b.sum() - 1) < 1e-2      # measured 1e-6
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
    """Directional derivative of loss(render(sigma_t + eps * dir)) by central differences with common random numbers
    (same seed, fixed majorant) vs <grad, dir> from render_backward.  Delta tracking makes a same-seed render piecewise
    constant in sigma_t per path (null/real decisions flip), so the FD of a finite-sample render carries
    O(sqrt(1/(eps*N))) noise: the check is a paired Z-test over independent seeds (|mean diff| < 3.5 standard errors)
    plus an absolute bound of 1.5 % on the mean.  The closed-form test below pins the gradient to 3e-3 without FD noise."""
    res, spp, gr, n_seeds = 16, 16384, 4, 6
    d = _small_inverse(res=res, spp=spp, grid_res=gr)
    dens0 = np.array(d["sphere"]["interior"]["sigma_t"]["grid"], dtype=np.float32)
    sc = mi.load_dict(d); ctx = mi.Context(0).upload(sc)
    rng = np.random.default_rng(12345)
    target = rng.random((res, res, 3)).astype(np.float32) * 0.5

    def loss_and_img(dens, seed):
        ctx.grid_update(0, np.ascontiguousarray(dens, dtype=np.float32))
        img = ctx.render(seed=seed, spp=spp)
        return float(((img.astype(np.float64) - target) ** 2).mean()), img

    dirs = [np.ones_like(dens0, dtype=np.float64)]
    zz = np.arange(gr)[:, None, None, None] / (gr - 1.0)
    dirs.append(np.broadcast_to(zz, dens0.shape).astype(np.float64))                 # linear ramp along z
    dirs.append(np.broadcast_to(1.0 - zz.transpose(1, 0, 2, 3), dens0.shape).astype(np.float64))   # ramp along y
    eps = 2e-3; fd = np.zeros((n_seeds, len(dirs))); an = np.zeros_like(fd)
    for s_i in range(n_seeds):
        seed = 100 + s_i
        _, img0 = loss_and_img(dens0, seed)
        grad_in = (2.0 * (img0.astype(np.float64) - target) / img0.size).astype(np.float32)
        ctx.grid_grad_zero(0)
        ctx.render_backward(grad_in, seed=seed, spp=spp)          # same seed => the adjoint replays the primal paths
        grad = ctx.grid_grad_read(0).reshape(dens0.shape).astype(np.float64)
        assert np.isfinite(grad).all() and np.abs(grad).max() > 0
        for k, m in enumerate(dirs):
            lp, _ = loss_and_img(dens0 + eps * m, seed); lm, _ = loss_and_img(dens0 - eps * m, seed)
            fd[s_i, k] = (lp - lm) / (2 * eps); an[s_i, k] = float((grad * m).sum())
    ctx.grid_update(0, dens0)
    diff = fd - an; mean = diff.mean(0); se = diff.std(0, ddof=1) / np.sqrt(n_seeds); ref = np.abs(an.mean(0))
    print("FD check: fd", fd.mean(0), "analytic", an.mean(0), "rel diff", mean / ref, "rel SE", se / ref)
    assert np.all(np.abs(mean) < 3.5 * se + 1e-3 * ref), (mean / ref, se / ref)
    assert np.all(np.abs(mean) / ref < 1.5e-2), mean / ref


def test_gradient_closed_form_absorbing_slab():
    """Purely absorbing uniform medium behind an index-matched (null) cube under a constant environment:
    pixel = exp(-scale * rho * chord) exactly, so d loss / d rho is known in closed form and the sum of the grid
    gradient must match it (no finite-difference noise; remaining error = Monte-Carlo noise of the adjoint, ~1e-3)."""
    res, spp, gr, scale, rho = 16, 16384, 4, 2.0, 0.35
    grid = np.full((gr, gr, gr, 1), rho, dtype=np.float32)
    d = {"type": "scene", "integrator": {"type": "prbvolpath", "max_depth": 8, "rr_depth": 64},
         "sensor": {"type": "perspective", "fov": 30.0, "near_clip": 0.01, "far_clip": 100.0,
                    "to_world": mi.T().look_at(origin=[0.3, 0.4, -4.0], target=[0, 0, 0], up=[0, 1, 0]),
                    "sampler": {"type": "independent", "sample_count": spp},
                    "film": {"type": "hdrfilm", "width": res, "height": res, "rfilter": {"type": "box"}, "pixel_format": "rgb"}},
         "cube": {"type": "cube", "to_world": mi.T().scale(0.5), "bsdf": {"type": "null"},
                  "interior": {"type": "heterogeneous", "scale": scale,
                               "sigma_t": {"type": "gridvolume", "grid": grid, "max_value": 0.5,
                                           "to_world": mi.T().translate([-0.5, -0.5, -0.5])},
                               "albedo": {"type": "rgb", "value": [0.0, 0.0, 0.0]}}},
         "env": {"type": "constant", "radiance": {"type": "rgb", "value": [1, 1, 1]}}}
    sc = mi.load_dict(d); ctx = mi.Context(0).upload(sc)
    img = ctx.render(seed=1, spp=spp).astype(np.float64)
    # closed form by quadrature: 16x16 stratified sub-samples per pixel, camera rays from the oracle's camera model,
    # chord through the cube [-.5,.5]^3 by the slab method:  I = <exp(-s rho l)>,  dI/drho = <-s l exp(-s rho l)>
    import ctypes as C
    o = orc.OracleScene(sc); ns = 16; ray = np.zeros(7, dtype=np.float32)
    I = np.zeros((res, res)); dI = np.zeros((res, res))
    for py in range(res):
        for px in range(res):
            acc = acc_d = 0.0
            for sy in range(ns):
                for sx in range(ns):
                    o.L.orc_kat_camera_ray(o.h, (px + (sx + .5) / ns) / res, (py + (sy + .5) / ns) / res, C.c_void_p(ray.ctypes.data))
                    ro, rd = ray[:3].astype(np.float64), ray[3:6].astype(np.float64)
                    with np.errstate(divide="ignore"):
                        t0 = (-0.5 - ro) / rd; t1 = (0.5 - ro) / rd
                    tn = np.minimum(t0, t1).max(); tf = np.maximum(t0, t1).min(); l = max(tf - max(tn, 0.0), 0.0)
                    e = np.exp(-scale * rho * l); acc += e; acc_d += -scale * l * e
            I[py, px] = acc / ns ** 2; dI[py, px] = acc_d / ns ** 2
    err = np.abs(img[..., 0] - I)                                    # primal: delta-tracking estimate of the transmittance (std <= 0.5/sqrt(spp))
    assert err.max() < 2e-2 and err.mean() < 2e-3, (err.max(), err.mean())
    target = np.zeros_like(img)
    grad_in = (2.0 * (img - target) / img.size).astype(np.float32)
    ctx.grid_grad_zero(0); ctx.render_backward(grad_in, seed=1, spp=spp)
    g = ctx.grid_grad_read(0).astype(np.float64).sum()
    closed = float((grad_in.astype(np.float64).sum(axis=2) * dI).sum())
    print("closed-form gradient check:", g, closed, abs(g - closed) / abs(closed))
    assert abs(g - closed) / abs(closed) < 1e-2        # Monte-Carlo noise of the adjoint at 4.2 M paths + quadrature error; measured 4e-3


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


def test_wavefront_prb_equals_per_lane_kernel(monkeypatch):
    """prbvolpath as a wavefront (primal pass + adjoint replay through queues) vs the per-lane kernel: same PCG32 streams,
    same draw order => lane-for-lane equal primal radiance and the same gradient up to fp32 summation order."""
    sc = mi.load_dict(_small_inverse(res=24, spp=64)); n = 24 * 24 * 64
    rng = np.random.default_rng(3); g = rng.standard_normal((24, 24, 3)).astype(np.float32)
    wf = mi.Context(0).upload(sc); a = wf.sample_lanes(5, 64, 0, n)
    wf.grid_grad_zero(0); wf.render_backward(g, seed=7, spp=64); ga = wf.grid_grad_read(0); st = wf.stats()
    monkeypatch.setenv("LR_MEGAKERNEL", "1")
    mk = mi.Context(0).upload(sc); b = mk.sample_lanes(5, 64, 0, n)
    mk.grid_grad_zero(0); mk.render_backward(g, seed=7, spp=64); gb = mk.grid_grad_read(0)
    assert st["launches"] > 16
    assert np.array_equal(a[:, 3], b[:, 3])
    frac = float(((np.abs(a[:, :3] - b[:, :3]) / np.maximum(np.abs(b[:, :3]), 1e-3)).max(axis=1) < 1e-5).mean())
    rel = float(np.linalg.norm(ga - gb) / np.linalg.norm(gb))
    print("wavefront vs per-lane prb: lanes equal", frac, "gradient rel L2", rel, "sum ratio", ga.sum() / gb.sum())
    assert frac > 0.999                     # measured 1.0; head-room for an ulp flipping a null/real decision
    assert rel < 1e-2 and abs(ga.sum() / gb.sum() - 1) < 5e-3      # measured 3e-7


def test_dda_adjoint_closed_form():
    """The closed-form absorbing-cube gradient with the production (majorant-DDA) tracking mode: the local majorants are
    detached like the reference's global one, the estimator stays unbiased."""
    res, spp, gr, scale, rho = 12, 16384, 8, 2.0, 0.35
    grid = np.full((gr, gr, gr, 1), rho, dtype=np.float32); grid[: gr // 2] *= 0.5           # two density levels => different local majorants
    def scene(tracking):
        return mi.load_dict({"type": "scene", "integrator": {"type": "prbvolpath", "max_depth": 8, "rr_depth": 64, "tracking": tracking, "majorant_cell": 2},
            "sensor": {"type": "perspective", "fov": 30.0, "near_clip": 0.01, "far_clip": 100.0,
                       "to_world": mi.T().look_at(origin=[0.3, 0.4, -4.0], target=[0, 0, 0], up=[0, 1, 0]),
                       "sampler": {"type": "independent", "sample_count": spp},
                       "film": {"type": "hdrfilm", "width": res, "height": res, "rfilter": {"type": "box"}, "pixel_format": "rgb"}},
            "cube": {"type": "cube", "to_world": mi.T().scale(0.5), "bsdf": {"type": "null"},
                     "interior": {"type": "heterogeneous", "scale": scale,
                                  "sigma_t": {"type": "gridvolume", "grid": grid, "max_value": 2 * rho,      # sigma_n > 0 everywhere: the detached
                                              "to_world": mi.T().translate([-0.5, -0.5, -0.5])},            # gradient estimator is singular at sigma_n = 0
                                  "albedo": {"type": "rgb", "value": [0.0, 0.0, 0.0]}}},
            "env": {"type": "constant", "radiance": {"type": "rgb", "value": [1, 1, 1]}}})
    out = {}
    for tr in ("reference", "dda"):
        ctx = mi.Context(0).upload(scene(tr)); img = ctx.render(seed=1, spp=spp).astype(np.float64)
        g_in = np.full((res, res, 3), 1.0 / (res * res * 3), dtype=np.float32)
        ctx.grid_grad_zero(0); ctx.render_backward(g_in, seed=1, spp=spp)
        out[tr] = (img.mean(), ctx.grid_grad_read(0).astype(np.float64).reshape(gr, gr, gr), ctx.stats()["collisions"])
    (m_r, g_r, c_r), (m_d, g_d, c_d) = out["reference"], out["dda"]
    assert abs(m_r - m_d) / m_r < 2e-3
    # per-z-slab sums of the gradient (128 voxels each) agree between the two estimators
    s_r, s_d = g_r.sum(axis=(1, 2)), g_d.sum(axis=(1, 2))
    assert np.all(np.abs(s_r - s_d) < 3e-2 * np.abs(s_r).max()), (s_r, s_d)
    assert abs(g_r.sum() - g_d.sum()) / abs(g_r.sum()) < 1e-2
    assert c_d < c_r


@pytest.mark.parametrize("tracking", ["reference", "dda"])
def test_albedo_grid_gradient(tracking):
    """The gradient w.r.t. an rgb ALBEDO gridvolume (row a24: prbvolpath.py:202-204 through sigma_s = sigma_t * albedo).  Reference order:
    CUDA adjoint == oracle adjoint on the same streams.  Both modes: the adjoint equals central finite differences of the CUDA primal with
    the same seed (the albedo never enters a sampling decision with roulette off, so the image is a polynomial of the albedo values: no
    Monte-Carlo noise in the differences), for a uniform shift per channel and for a single voxel; sigma_t and albedo gradients coexist."""
    from tests.test_oracle_golden import _albedo_grid_scene
    res, spp = 16, 64
    sc = mi.load_dict(_albedo_grid_scene(res, spp, tracking)); ctx = mi.Context(0).upload(sc)
    ga = sc.param_keys["cube.interior_medium.albedo.data"]; gs = sc.param_keys["cube.interior_medium.sigma_t.data"]
    rng = np.random.default_rng(9); w = rng.random((res, res, 3)).astype(np.float32)
    ctx.grid_grad_zero(ga); ctx.grid_grad_zero(gs); ctx.render_backward(w, seed=21, spp=spp)
    grad = ctx.grid_grad_read(ga).reshape(4, 4, 4, 3).astype(np.float64); gsig = ctx.grid_grad_read(gs)
    assert np.isfinite(grad).all() and np.abs(grad).max() > 0 and np.abs(gsig).max() > 0
    if tracking == "reference":
        o = orc.OracleScene(sc); o.grid_grad_zero(ga); o.grid_grad_zero(gs); o.render_backward(w, seed=21, spp=spp)
        b = o.grid_grad_read(ga).reshape(4, 4, 4, 3)
        assert np.linalg.norm(grad - b) / np.linalg.norm(b) < 2e-2 and abs(grad.sum() / b.sum() - 1) < 1e-2
        assert np.linalg.norm(gsig - o.grid_grad_read(gs)) / np.linalg.norm(o.grid_grad_read(gs)) < 2e-2
    base = np.array(sc.grid_arrays[ga], dtype=np.float32)
    def loss(a):
        ctx.grid_update(ga, a); return float((ctx.render(seed=21, spp=spp).astype(np.float64) * w).sum())
    eps = 2e-2
    for c in range(3):
        d = np.zeros_like(base); d[..., c] = eps
        fd = (loss(base + d) - loss(base - d)) / (2 * eps)
        assert abs(fd - grad[..., c].sum()) <= 3e-3 * abs(fd) + 1e-5, (c, fd, grad[..., c].sum())
    d = np.zeros_like(base); d[1, 2, 1, 0] = eps
    fd = (loss(base + d) - loss(base - d)) / (2 * eps)
    assert abs(fd - grad[1, 2, 1, 0]) <= 1e-2 * abs(fd) + 1e-5, (fd, grad[1, 2, 1, 0])
    # a 1-channel albedo grid receives the sum over the channels
    d1 = _albedo_grid_scene(res, spp, tracking); d1["cube"]["interior"]["albedo"]["grid"] = np.ascontiguousarray(base[..., :1])
    sc1 = mi.load_dict(d1); c1 = mi.Context(0).upload(sc1); c1.grid_grad_zero(1); c1.render_backward(w, seed=21, spp=spp); g1 = c1.grid_grad_read(1).astype(np.float64)
    def loss1(a):
        c1.grid_update(1, a); return float((c1.render(seed=21, spp=spp).astype(np.float64) * w).sum())
    b1 = np.ascontiguousarray(base[..., :1]); fd = (loss1(b1 + eps) - loss1(b1 - eps)) / (2 * eps)
    assert abs(fd - g1.sum()) <= 3e-3 * abs(fd) + 1e-5, (fd, g1.sum())


def test_prb_track_finishes_sealed_scatter_vertices_itself(monkeypatch):
    """prbvolpath in production tracking with a sealed medium: TRACK shades its scatter vertices itself (wvp_track_scatter_sealed).  Same draws
    in the same order as the three-kernel form (LR_TRACK_FUSE_OFF=1): identical primal lanes, the same gradient up to fp32 atomic order."""
    sc = mi.load_dict(scenes.sphere_liver_inverse(res=48, spp=8, grid_res=32, tracking="dda")); n = 48 * 48 * 8
    out = {}
    for key, env in (("fused", None), ("plain", "1")):
        if env is None:
            monkeypatch.delenv("LR_TRACK_FUSE_OFF", raising=False)
        else:
            monkeypatch.setenv("LR_TRACK_FUSE_OFF", env)
        ctx = mi.Context(0).upload(sc)
        lanes = ctx.sample_lanes(2, 8, 0, n); img = ctx.render(seed=2, spp=8)
        g = np.ones_like(img) / img.size
        ctx.grid_grad_zero(0); ctx.render_backward(g, seed=3, spp=8); grad = ctx.grid_grad_read(0).copy()
        out[key] = (lanes, grad, ctx.stats())
    assert np.array_equal(out["fused"][0], out["plain"][0])
    ga, gb = out["fused"][1], out["plain"][1]
    assert np.abs(gb).max() > 0
    np.testing.assert_allclose(ga, gb, rtol=2e-4, atol=1e-7 * np.abs(gb).max())
    assert out["fused"][2]["collisions"] == out["plain"][2]["collisions"]
