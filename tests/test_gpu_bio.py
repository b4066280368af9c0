"""GPU parity for SURVEY.md 8(f) rank 2: the fork's tissue media (liver / glissonCapsule / parenchyma) and its biovolpath
integrator on the CUDA path (csrc/lr_biovolpath.cuh), called through the C ABI, against the oracle's restatement
(oracle/orc_biovolpath.inl) on the same seeded inputs.  Tolerances: the free-path sampler agrees to the last ulps of
logf / log2f (device vs libm); per-path radiance within 1e-3 relative on >= 97 % of the lanes (identical PCG32 streams =>
identical paths until an ulp flips a branch), image means within 2 %."""
import numpy as np
import pytest

import liverrenderer_b200 as mi
from liverrenderer_b200 import _capi as K
from liverrenderer_b200 import scenes
from tests import orc
from tests.test_gpu_parity import path_match_fraction
from tests.test_oracle_bio import _medium

pytestmark = pytest.mark.gpu
F = np.float32


@pytest.mark.parametrize("kind", ["liver", "glissonCapsule", "parenchyma"])
def test_tissue_free_path_sampler_matches_oracle(kind):
    m = _medium(kind); rng = np.random.default_rng(11); n = 200_000
    sample = rng.random(n, dtype=F)
    depth = np.where(rng.random(n) < 0.5, rng.random(n) * 0.012, rng.random(n) * 2).astype(F)
    maxt = np.where(rng.random(n) < 0.3, np.inf, rng.random(n) * 3).astype(F)
    ch = rng.integers(0, 3, n).astype(np.uint32); act = (rng.random(n) < 0.9).astype(np.uint32)
    sample[:3] = [0.0, np.nextafter(F(1), F(0)), 0.5]
    ctx = mi.Context(0)
    gpu = ctx.kat_tissue(m, sample, depth, maxt, ch, act); cpu = orc.tissue_probe(m, sample, depth, maxt, ch, act)
    fin = np.isfinite(cpu[:, 0])
    assert np.array_equal(np.isfinite(gpu[:, 0]), fin)
    rel = np.abs(gpu[fin, 0] - cpu[fin, 0]) / np.abs(cpu[fin, 0])
    assert rel.max() < 1e-6, rel.max()                                     # a few ulps of logf / log2f
    assert (rel == 0).mean() > 0.5
    # element type, transmittance, t, coefficients: equal except where an ulp of the distance flips a comparison
    same = np.all((gpu[:, 1:] == cpu[:, 1:]) | (np.isclose(gpu[:, 1:], cpu[:, 1:], rtol=1e-6, atol=0)), axis=1)
    assert same.mean() > 0.9995, same.mean()


def _rough(**kw):
    """Liver-SingleMesh with a rough dielectric boundary: a smooth lobe, so emitter sampling and the shadow walk through tissue run
    (the fork's own boundaries are delta lobes: no emitter sampling anywhere in its scenes)."""
    return scenes.liver_single_mesh_bio(surface="roughdielectric", **kw)


@pytest.mark.parametrize("builder,sampler", [(scenes.liver_single_mesh_bio, "independent"), (scenes.liver_multi_mesh_bio, "ldsampler"),
                                             (scenes.liver_multi_mesh_bio, "independent"), (_rough, "independent")])
def test_biovolpath_per_path(builder, sampler):
    res, spp = 48, 16
    sc = mi.load_dict(builder(res=res, spp=spp, sampler=sampler)); n = res * res * spp
    ctx = mi.Context(0).upload(sc)
    gpu = ctx.sample_lanes(0, spp, 0, n); o = orc.OracleScene(sc); cpu = o.sample_lanes(0, spp, 0, n)
    assert np.isfinite(gpu).all() and np.array_equal(gpu[:, 3], cpu[:, 3])
    frac = path_match_fraction(gpu, cpu, 1e-3)
    assert frac > 0.97, frac
    m_g, m_c = gpu[:, :3].mean(0), cpu[:, :3].mean(0)
    assert np.all(np.abs(m_g - m_c) / m_c < 2e-2), (m_g, m_c)
    st, so = ctx.stats(), o.stats()
    assert abs(st["real_scatters"] - so["real_scatters"]) / so["real_scatters"] < 3e-2
    # the result-clearing rule of biovolpath.cpp:304 on the device: three live channels, one, or none -- never two
    live = (gpu[:, :3] > 0).sum(1)
    assert (live == 2).sum() == 0 and (live == 1).sum() > 0 and (live == 0).sum() > 0


def test_biovolpath_film_matches_oracle_film():
    res, spp = 32, 32
    sc = mi.load_dict(scenes.liver_single_mesh_bio(res=res, spp=spp))
    img = mi.Context(0).upload(sc).render(seed=3, spp=spp)
    o = orc.OracleScene(sc); o.render(3, spp); ref = o.film_develop()
    assert img.shape == ref.shape and np.isfinite(img).all()
    np.testing.assert_allclose(img[..., 3], ref[..., 3], rtol=1e-5)
    # same lanes, same splats: pixels differ only through the < 3 % of lanes an ulp sends down another path
    close = np.isclose(img[..., :3], ref[..., :3], rtol=1e-3, atol=1e-5).all(-1)
    assert close.mean() > 0.7, close.mean()
    assert abs(img[..., :3].mean() - ref[..., :3].mean()) / ref[..., :3].mean() < 2e-2


def test_biovolpath_max_depth_and_roulette_edges():
    """The dying-iteration rule (lr_biovolpath.cuh note (2)) is exercised hardest by short paths: max_depth 1, 2, 3."""
    for md in (1, 2, 3):
        sc = mi.load_dict(scenes.liver_single_mesh_bio(res=32, spp=8, max_depth=md)); n = 32 * 32 * 8
        gpu = mi.Context(0).upload(sc).sample_lanes(0, 8, 0, n); cpu = orc.OracleScene(sc).sample_lanes(0, 8, 0, n)
        assert path_match_fraction(gpu, cpu, 1e-3) > 0.985, md
        assert np.array_equal((gpu[:, :3] == 0).all(1), (cpu[:, :3] == 0).all(1)) or \
            ((gpu[:, :3] == 0).all(1) != (cpu[:, :3] == 0).all(1)).mean() < 0.01


def test_tissue_media_and_biovolpath_come_together():
    """Medium::sample_interaction with a tissue depth throws for every other medium (src/render/medium.cpp:84-90)."""
    d = scenes.liver_single_mesh_bio(res=8, spp=1); d["integrator"]["type"] = "volpath"
    with pytest.raises(RuntimeError, match="only sampled by biovolpath"):
        mi.Context(0).upload(mi.load_dict(d))
    d = scenes.liver_single_mesh(res=8, spp=1, grid_res=8); d["integrator"] = {"type": "biovolpath", "max_depth": 4}
    with pytest.raises(RuntimeError, match="biovolpath needs liver"):
        mi.Context(0).upload(mi.load_dict(d))
    ctx = mi.Context(0)
    with pytest.raises(RuntimeError, match="not a tissue medium"):
        m = _medium("liver"); m.kind = K.MEDIUM_HOMOGENEOUS
        ctx.kat_tissue(m, [0.5], 0.0, 1.0, 0)


def test_biovolpath_lane_partition_is_exact():
    sc = mi.load_dict(scenes.liver_single_mesh_bio(res=16, spp=4)); n = 16 * 16 * 4
    ctx = mi.Context(0).upload(sc)
    whole = ctx.sample_lanes(0, 4, 0, n)
    parts = np.concatenate([ctx.sample_lanes(0, 4, 0, 300), ctx.sample_lanes(0, 4, 300, n)])
    assert np.array_equal(whole, parts)


@pytest.mark.parametrize("builder", [scenes.liver_single_mesh_bio, scenes.liver_multi_mesh_bio, _rough])
def test_biovolpath_wavefront_equals_per_lane_kernel(builder, monkeypatch):
    """The production path (lr_wavefront_bio.cuh: stage kernels over queues, TRACE stage) and k_render<biovolpath> draw the same
    PCG32 streams in the same order: lane-for-lane equal radiance, equal event counts."""
    res, spp = 40, 16
    sc = mi.load_dict(builder(res=res, spp=spp)); n = res * res * spp
    wf = mi.Context(0).upload(sc); a = wf.sample_lanes(3, spp, 0, n); st_a = wf.stats()
    monkeypatch.setenv("LR_BIO_PER_LANE", "1")
    pl = mi.Context(0).upload(sc); b = pl.sample_lanes(3, spp, 0, n); st_b = pl.stats()
    assert st_a["launches"] > 8 and st_b["launches"] <= 2           # really two different code paths
    assert np.array_equal(a[:, 3], b[:, 3])
    from tests.test_gpu_parity import rel_err
    assert (rel_err(a[:, :3], b[:, :3], 1e-3).max(axis=1) < 1e-5).mean() > 0.9995
    assert st_a["real_scatters"] == st_b["real_scatters"] and st_a["collisions"] == st_b["collisions"]
    # and the film of the two agrees (atomics in another order)
    monkeypatch.delenv("LR_BIO_PER_LANE")
    img_w = mi.Context(0).upload(sc).render(seed=5, spp=spp)
    monkeypatch.setenv("LR_BIO_PER_LANE", "1")
    img_p = mi.Context(0).upload(sc).render(seed=5, spp=spp)
    np.testing.assert_allclose(img_w, img_p, rtol=2e-4, atol=1e-6)


def test_biovolpath_wavefront_small_pool_multi_batch(monkeypatch):
    """More lanes than pool slots: several lock-step batches on several streams give the same film."""
    sc = mi.load_dict(scenes.liver_single_mesh_bio(res=64, spp=16))
    big = mi.Context(0).upload(sc).render(seed=2, spp=16)
    monkeypatch.setenv("LR_WF_SLOTS", "8192")
    small = mi.Context(0).upload(sc).render(seed=2, spp=16)
    assert np.allclose(big, small, rtol=2e-4, atol=1e-6)


def test_biovolpath_sample_index_split_adds_up():
    """SURVEY 8(e) for this row: the frame split by sample index (lr_set_lane_map) -- two half-renders accumulate to the whole one."""
    res, spp = 32, 16
    sc = mi.load_dict(scenes.liver_single_mesh_bio(res=res, spp=spp))
    ctx = mi.Context(0).upload(sc)
    ctx.film_clear(); ctx.render_lanes(4, spp, 0, 0); whole = ctx.film_read()
    ctx.film_clear()
    for r in range(2):
        ctx.set_lane_map(spp, r * spp // 2, spp // 2); ctx.render_lanes(4, spp, 0, 0)
    parts = ctx.film_read(); ctx.set_lane_map(0, 0, 0)
    np.testing.assert_allclose(parts[..., -1], whole[..., -1], rtol=0, atol=0)          # filter weights: exact
    np.testing.assert_allclose(parts[..., :3], whole[..., :3], rtol=2e-5, atol=1e-6)    # radiance: fp32 atomics in another order
