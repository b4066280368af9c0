import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes
from tests import orc
res, spp = 64, 256
def scene(tr):
    d = scenes.liver_multi_mesh(res=res, spp=spp, tracking=tr)
    d["capsule"]["sigma_t"]["grid"] = scenes.sigmat_field(64); d["parenchyma"]["sigma_t"]["grid"] = scenes.sigmat_field(128)
    return d
n = res * res * spp
for tr in ("reference", "dda"):
    ctx = mi.Context(0).upload(mi.load_dict(scene(tr))); a = ctx.sample_lanes(11, spp, 0, n); bad = ~np.isfinite(a).all(1)
    print("gpu", tr, "non-finite lanes", bad.sum(), np.nonzero(bad)[0][:5], a[bad][:3])
o = orc.OracleScene(mi.load_dict(scene("reference"))); c = o.sample_lanes(5, spp, 0, n); bad = ~np.isfinite(c).all(1)
print("oracle non-finite lanes", bad.sum(), np.nonzero(bad)[0][:5], c[bad][:3])
