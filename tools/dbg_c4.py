import os, sys, copy
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes
res, spp = 48, 2048
n = res * res * spp
def scene(tr, md):
    d = scenes.liver_multi_mesh(res=res, spp=spp, tracking=tr)
    d["capsule"]["sigma_t"]["grid"] = np.full((8, 8, 8, 1), 2.5, np.float32); d["parenchyma"]["sigma_t"]["grid"] = np.full((8, 8, 8, 1), 2.5, np.float32)
    d["outer"]["bsdf"] = {"type": "roughdielectric", "int_ior": 1.38, "ext_ior": 1.0, "alpha": 0.1}
    d["integrator"]["max_depth"] = md
    return d
for env in ({}, {"LR_SHADOW_CLOSEST": "1"}, {"LR_NO_SEAL": "1", "LR_SHADOW_CLOSEST": "1"}):
    for k, v in env.items(): os.environ[k] = v
    r = {}
    for tr in ("reference", "dda"):
        ctx = mi.Context(0).upload(mi.load_dict(scene(tr, 4)))
        a = ctx.sample_lanes(3, spp, 0, n)[:, :3].astype(np.float64); st = ctx.stats(); ctx.close(); r[tr] = (a.mean(0), st["rays"] / st["samples"])
    for k in env: os.environ.pop(k)
    print(env, "ref", r["reference"], "dda", r["dda"])
