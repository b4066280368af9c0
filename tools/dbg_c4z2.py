"""C4: GPU-DDA vs GPU-reference-order at growing spp (does the per-pixel z spread shrink to 1 = heavy tails, or grow = bias?)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes
res = 64
def scene(tr, spp):
    d = scenes.liver_multi_mesh(res=res, spp=spp, tracking=tr)
    d["capsule"]["sigma_t"]["grid"] = scenes.sigmat_field(64); d["parenchyma"]["sigma_t"]["grid"] = scenes.sigmat_field(128)
    return d
def moments(tr, seed, spp, rows=4):
    ctx = mi.Context(0).upload(mi.load_dict(scene(tr, spp))); m = np.zeros((res * res, 3)); v = np.zeros((res * res, 3)); zf = np.zeros(res * res)
    for y in range(0, res, rows):
        b, e = y * res * spp, (y + rows) * res * spp
        a = ctx.sample_lanes(seed, spp, b, e)[:, :3].reshape(rows * res, spp, 3).astype(np.float64)
        m[y * res:(y + rows) * res] = a.mean(1); v[y * res:(y + rows) * res] = a.var(1, ddof=1) / spp; zf[y * res:(y + rows) * res] = (a.sum(2) == 0).mean(1)
    ctx.close(); return m, v, zf
for spp in (4096, 32768):
    ma, va, zfa = moments("dda", 11, spp); mb, vb, zfb = moments("reference", 13, spp)
    se = np.sqrt(va + vb); ok = se > 0; z = (ma - mb)[ok] / se[ok]
    print(f"spp {spp}: frac3 {(np.abs(z) < 3).mean():.4f} mean z {z.mean():+.3f} rms z {np.sqrt((z**2).mean()):.3f}")
    organ = zfb > 0.3
    for name, sel in (("organ (zero-frac>0.3)", organ), ("rest", ~organ)):
        d = (ma[sel].sum(0) - mb[sel].sum(0)); s = np.sqrt((va[sel] + vb[sel]).sum(0)); base = mb[sel].sum(0)
        print(f"   {name}: {sel.sum()} px, region bias {(d / base).round(5)} +- {(s / base).round(5)}")
