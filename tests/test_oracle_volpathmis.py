"""SURVEY.md 8(f) rank 4, first half: volpathmis (src/integrators/volpathmis.cpp) in the oracle (oracle/orc_volpathmis.inl).
PINNED by the reference's own golden images: INTEGRATOR_MAPPING (src/render/tests/test_renders.py:45-49) renders every
`*_volpath` scene with volpathmis as well and Z-tests it against the same reference image -- with the default
use_spectral_mis = true, which is therefore the pinned configuration."""
import numpy as np
import pytest

import liverrenderer_b200 as mi
from liverrenderer_b200 import _capi as K
from tests import orc, golden_scenes as G

SPP = int(2e6) // (48 * 48)


def _scene(name, spectral=True, spp=SPP, **kw):
    d = G.scene(name, integrator="volpathmis", spp=spp, **kw); d["integrator"]["use_spectral_mis"] = spectral
    return mi.load_dict(d)


@pytest.mark.parametrize("name", ["test_hetero_sigmat_albedo", "test_homo_dielectric_envmap"])
def test_oracle_volpathmis_passes_reference_ztest(name):
    sc = _scene(name)
    assert sc.integrator.kind == K.INTEGRATOR_VOLPATHMIS and sc.integrator.spectral_mis == 1
    img = orc.OracleScene(sc).render(seed=0, spp=SPP)
    frac, ok = G.reference_accepts(img, name, SPP)
    assert ok, f"{name} (volpathmis): only {frac:.4f} accepted"
    assert np.isfinite(img).all()          # p / f bookkeeping maps every non-finite ratio to 0 (volpathmis.cpp:641-651)


def test_plain_mis_on_a_grey_medium_passes_the_ztest_and_loses_energy_on_a_chromatic_one():
    """use_spectral_mis = false keeps one row of p / f and declares a path dead as soon as ANY channel of it is 0
    (mis_weight, volpathmis.cpp:665-667): unbiased for a grey medium, darker where the albedo grid has pure colours
    (31 % of the golden albedo texels have a zero channel).  The reference's tests only run the default."""
    name = "test_hetero_sigmat"
    img = orc.OracleScene(_scene(name, spectral=False)).render(seed=0, spp=SPP)
    assert G.reference_accepts(img, name, SPP)[1]
    name = "test_hetero_sigmat_albedo"; spp = 128
    a = orc.OracleScene(_scene(name, spectral=True, spp=spp)).render(seed=0, spp=spp)[..., :3].mean()
    b = orc.OracleScene(_scene(name, spectral=False, spp=spp)).render(seed=0, spp=spp)[..., :3].mean()
    assert b < a * 0.995


def test_volpathmis_agrees_with_volpath_on_a_surface_scene():
    """Same estimand, different MIS heuristic (balance over p / f sums, volpathmis.cpp:672-688, instead of volpath's power
    heuristic): the images agree statistically."""
    from liverrenderer_b200 import scenes
    d = scenes.cornell_box_path(res=32, spp=256); d["integrator"] = {"type": "volpath", "max_depth": 6}
    a = orc.OracleScene(mi.load_dict(d)).render(seed=0, spp=256)[..., :3]
    d["integrator"] = {"type": "volpathmis", "max_depth": 6}
    b = orc.OracleScene(mi.load_dict(d)).render(seed=1, spp=256)[..., :3]
    assert abs(a.mean() - b.mean()) / a.mean() < 0.02
    blk = lambda x: x.reshape(4, 8, 4, 8, 3).mean((1, 3))
    np.testing.assert_allclose(blk(b), blk(a), rtol=0.12, atol=0.02)
