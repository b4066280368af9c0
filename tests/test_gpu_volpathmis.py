"""GPU parity for SURVEY.md 8(f) rank 4 (first half): volpathmis on the CUDA path (csrc/lr_volpathmis.cuh) through the C ABI,
(1) lane for lane against the oracle's restatement on the same PCG32 streams and (2) against the reference's own golden
images with the reference's acceptance test (test_renders.py renders every *_volpath scene with volpathmis too)."""
import numpy as np
import pytest

import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes
from tests import orc, golden_scenes as G
from tests.test_gpu_parity import path_match_fraction
from tests.test_oracle_volpathmis import _scene, SPP

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", G.GOLDEN_SCENES)
def test_gpu_volpathmis_passes_reference_ztest(name):
    sc = _scene(name)
    img = mi.Context(0).upload(sc).render(seed=0, spp=SPP)
    frac, ok = G.reference_accepts(img, name, SPP)
    assert ok, f"{name} (volpathmis, GPU): only {frac:.4f} accepted"
    assert np.isfinite(img).all()


@pytest.mark.parametrize("name,spectral", [("test_hetero_sigmat_albedo", True), ("test_hetero_sigmat_albedo", False),
                                           ("test_homo_null_envmap", True), ("test_homo_dielectric_envmap", False)])
def test_gpu_volpathmis_per_path(name, spectral):
    spp = 16; sc = _scene(name, spectral=spectral, spp=spp); n = 48 * 48 * spp
    ctx = mi.Context(0).upload(sc)
    gpu = ctx.sample_lanes(0, spp, 0, n); o = orc.OracleScene(sc); cpu = o.sample_lanes(0, spp, 0, n)
    assert np.array_equal(gpu[:, 3], cpu[:, 3])
    frac = path_match_fraction(gpu, cpu, 1e-3)
    assert frac > 0.97, frac
    m_g, m_c = gpu[:, :3].mean(0), cpu[:, :3].mean(0)
    assert np.all(np.abs(m_g - m_c) / m_c < 2e-2), (m_g, m_c)
    st, so = ctx.stats(), o.stats()
    assert abs(st["collisions"] - so["collisions"]) / max(so["collisions"], 1) < 5e-2


@pytest.mark.parametrize("spectral", [True, False])
def test_gpu_volpathmis_liver_scene_per_path(spectral):
    """The C3 plugin set (rough dielectric boundary, heterogeneous gridvolume, chromatic albedo, envmap) under volpathmis."""
    d = scenes.liver_single_mesh(res=40, spp=8, grid_res=32, phase="hg")
    d["integrator"] = {"type": "volpathmis", "max_depth": 12, "rr_depth": 5, "use_spectral_mis": spectral}
    sc = mi.load_dict(d); n = 40 * 40 * 8
    gpu = mi.Context(0).upload(sc).sample_lanes(0, 8, 0, n); cpu = orc.OracleScene(sc).sample_lanes(0, 8, 0, n)
    assert path_match_fraction(gpu, cpu, 1e-3) > 0.96
    m_g, m_c = gpu[:, :3].mean(0), cpu[:, :3].mean(0)
    assert np.all(np.abs(m_g - m_c) / m_c < 3e-2), (m_g, m_c)


def test_gpu_volpathmis_agrees_with_volpath_image():
    """Same estimand as volpath (the production wavefront): block means of the two images agree."""
    res, spp = 32, 256
    d = scenes.liver_single_mesh(res=res, spp=spp, grid_res=32)
    a = mi.Context(0).upload(mi.load_dict(d)).render(seed=0, spp=spp)[..., :3]
    d["integrator"] = {"type": "volpathmis", "max_depth": 12, "rr_depth": 5}
    b = mi.Context(0).upload(mi.load_dict(d)).render(seed=1, spp=spp)[..., :3]
    assert abs(a.mean() - b.mean()) / a.mean() < 0.02
    blk = lambda x: x.reshape(4, 8, 4, 8, 3).mean((1, 3))
    np.testing.assert_allclose(blk(b), blk(a), rtol=0.1, atol=0.02)


@pytest.mark.parametrize("spectral", [True, False])
def test_volpathmis_wavefront_equals_per_lane_kernel(spectral, monkeypatch):
    """The production path (lr_wavefront_mis.cuh: stage kernels over queues, TRACE stage, p / f state in its own records) and
    k_render<volpathmis> draw the same PCG32 streams in the same order: lane-for-lane equal radiance, equal event counts."""
    from tests.test_gpu_parity import rel_err
    d = scenes.liver_single_mesh(res=40, spp=8, grid_res=32, phase="hg")
    d["integrator"] = {"type": "volpathmis", "max_depth": 12, "rr_depth": 5, "use_spectral_mis": spectral}
    sc = mi.load_dict(d); n = 40 * 40 * 8
    wf = mi.Context(0).upload(sc); a = wf.sample_lanes(3, 8, 0, n); st_a = wf.stats()
    monkeypatch.setenv("LR_MIS_PER_LANE", "1")
    pl = mi.Context(0).upload(sc); b = pl.sample_lanes(3, 8, 0, n); st_b = pl.stats()
    assert st_a["launches"] > 8 and st_b["launches"] <= 2
    assert np.array_equal(a[:, 3], b[:, 3])
    assert (rel_err(a[:, :3], b[:, :3], 1e-3).max(axis=1) < 1e-5).mean() > 0.999
    assert st_a["collisions"] == st_b["collisions"] and st_a["real_scatters"] == st_b["real_scatters"]


def test_volpathmis_wavefront_golden_scene_equals_per_lane(monkeypatch):
    """Null boundaries, a homogeneous medium and hide-able area emitters (the golden scenes' plugin set) through both paths."""
    from tests.test_gpu_parity import rel_err
    for name in ("test_homo_null_envmap", "test_hetero_sigmat_albedo"):
        sc = _scene(name, spp=16); n = 48 * 48 * 16
        monkeypatch.delenv("LR_MIS_PER_LANE", raising=False)
        a = mi.Context(0).upload(sc).sample_lanes(1, 16, 0, n)
        monkeypatch.setenv("LR_MIS_PER_LANE", "1")
        b = mi.Context(0).upload(sc).sample_lanes(1, 16, 0, n)
        assert (rel_err(a[:, :3], b[:, :3], 1e-3).max(axis=1) < 1e-5).mean() > 0.999, name
