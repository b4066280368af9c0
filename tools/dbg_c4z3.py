"""C4 converged image, GPU-DDA vs oracle: (1) Russian roulette off (rr_depth > max_depth), (2) on, split by how heavy-tailed the oracle's pixel estimator is."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes
from tests import orc
res, spp = 64, 4096
n = res * res * spp
def scene(tr, **kw):
    d = scenes.liver_multi_mesh(res=res, spp=spp, tracking=tr)
    d["capsule"]["sigma_t"]["grid"] = scenes.sigmat_field(64); d["parenchyma"]["sigma_t"]["grid"] = scenes.sigmat_field(128)
    d["integrator"].update(kw); return d
for kw in ({"rr_depth": 1000}, {}):
    ctx = mi.Context(0).upload(mi.load_dict(scene("dda", **kw))); a = ctx.sample_lanes(11, spp, 0, n)[:, :3].reshape(res * res, spp, 3).astype(np.float64); ctx.close()
    o = orc.OracleScene(mi.load_dict(scene("reference", **kw))); b = o.sample_lanes(5, spp, 0, n)[:, :3].reshape(res * res, spp, 3).astype(np.float64)
    bad = ~np.isfinite(b).all(axis=2); b[bad] = 0
    ma, mb = a.mean(1), b.mean(1); va, vb = a.var(1, ddof=1) / spp, b.var(1, ddof=1) / spp
    se = np.sqrt(va + vb); ok = se > 0; z = np.where(ok, (ma - mb) / np.where(ok, se, 1), 0)
    mean = mb.mean(); mse = ((ma - mb) ** 2).mean(); mc = (va + vb).mean(); se_resid = np.sqrt(2 * ((va + vb) ** 2).mean() / ma.size)
    print(kw, f"frac3 {(np.abs(z[ok]) < 3).mean():.4f} mean z {z[ok].mean():+.3f} rms z {np.sqrt((z[ok]**2).mean()):.3f} non-finite {bad.sum()} resid {(mse - mc) / mean**2:.2e} (1e-6 allowed + 3se {3 * se_resid / mean**2:.2e}) bias {(ma.mean() - mb.mean()) / mean:+.2e} max orc {b.max():.1f} max dda {a.max():.1f}")
    w = (b ** 2).max(1) / np.maximum((b ** 2).sum(1), 1e-300)            # share of the largest sample in the second moment
    wa = (a ** 2).max(1) / np.maximum((a ** 2).sum(1), 1e-300)
    for thr in (0.01, 0.02, 0.05, 0.1):
        sel = ok & (w < thr) & (wa < thr)
        print(f"   max-sample share < {thr}: {sel.mean() * 100:.1f} % of pixel-channels, frac3 {(np.abs(z[sel]) < 3).mean():.4f} rms z {np.sqrt((z[sel]**2).mean()):.3f};  rest: frac3 {(np.abs(z[ok & ~sel]) < 3).mean():.4f} mean z {z[ok & ~sel].mean():+.3f}")
