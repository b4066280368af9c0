"""Which pixels of the C4 converged-image check disagree: z maps of GPU-DDA / GPU-reference-order / oracle against each other."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes
from tests import orc
res, spp = 64, 4096
n = res * res * spp
def scene(tr):
    d = scenes.liver_multi_mesh(res=res, spp=spp, tracking=tr)
    d["capsule"]["sigma_t"]["grid"] = scenes.sigmat_field(64); d["parenchyma"]["sigma_t"]["grid"] = scenes.sigmat_field(128)
    return d
def gpu(tr, seed):
    ctx = mi.Context(0).upload(mi.load_dict(scene(tr))); a = ctx.sample_lanes(seed, spp, 0, n)[:, :3].reshape(res * res, spp, 3).astype(np.float64); ctx.close(); return a
A = {"dda": gpu("dda", 11), "dda2": gpu("dda", 12), "ref": gpu("reference", 13)}
o = orc.OracleScene(mi.load_dict(scene("reference"))); c = o.sample_lanes(5, spp, 0, n)[:, :3].reshape(res * res, spp, 3).astype(np.float64)
bad = ~np.isfinite(c).all(axis=2); c[bad] = 0; A["orc"] = c
print("oracle non-finite", bad.sum(), "pixels", np.nonzero(bad.any(1))[0])
def cmp(x, y):
    a, b = A[x], A[y]
    ma, mb = a.mean(1), b.mean(1); va, vb = a.var(1, ddof=1) / spp, b.var(1, ddof=1) / spp
    se = np.sqrt(va + vb); ok = se > 0; z = np.where(ok, (ma - mb) / np.where(ok, se, 1), 0)
    print(f"{x} vs {y}: frac3 {(np.abs(z[ok]) < 3).mean():.4f}  mean z {z[ok].mean():+.3f}  rms z {np.sqrt((z[ok]**2).mean()):.3f}")
    return z
cmp("dda", "dda2"); cmp("dda", "ref"); cmp("ref", "orc"); z = cmp("dda", "orc")
out = np.nonzero((np.abs(z) >= 3).any(1))[0]
print("outlier pixels", len(out))
zm = np.abs(z).max(1).reshape(res, res)
for y in range(res):
    print("".join("#" if zm[y, x] >= 4 else ("+" if zm[y, x] >= 3 else ".") for x in range(res)))
for p in out[:40]:
    a, b = A["dda"][p], A["orc"][p]; r = A["ref"][p]
    print(f"px ({p % res},{p // res}) z {z[p].round(1)} dda mean {a.mean(0).round(4)} orc {b.mean(0).round(4)} gpuref {r.mean(0).round(4)} max dda {a.max():.1f} orc {b.max():.1f} zero-frac dda {(a.sum(1)==0).mean():.3f} orc {(b.sum(1)==0).mean():.3f}")
