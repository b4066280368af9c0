import sys, time, os
sys.path.insert(0, "/root/repo")
import numpy as np
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes
res, spp = 512, 16
def run(label, d):
    ctx = mi.Context(0).upload(mi.load_dict(d)); ctx.render(seed=0, spp=1)
    t = time.perf_counter(); ctx.render(seed=1, spp=spp); dt = time.perf_counter() - t
    st = ctx.stats()
    print(label, round(dt * 1e3, 1), "ms", round(res * res * spp / dt / 1e6, 2), "M/s", "coll/sample", round(st["collisions"] / st["samples"], 1),
          "Mcoll/s", round(st["collisions"] / dt / 1e6, 1), "rays/sample", round(st["rays"] / st["samples"], 2), flush=True)
mis = {"type": "volpathmis", "max_depth": 12, "rr_depth": 5}
for g in (32, 256):
    d = scenes.liver_single_mesh(res=res, spp=spp, grid_res=g, tracking="reference"); d["integrator"] = mis
    run(f"mis grid {g}", d)
d = scenes.liver_single_mesh(res=res, spp=spp, grid_res=256, tracking="reference"); d["integrator"] = mis; d["liver"]["interior"]["sample_emitters"] = False
run("mis grid 256 no medium NEE", d)
d = scenes.liver_single_mesh(res=res, spp=spp, grid_res=256, tracking="reference"); d["integrator"] = dict(mis, max_depth=2)
run("mis grid 256 max_depth 2", d)
d = scenes.liver_single_mesh(res=res, spp=spp, grid_res=256, tracking="reference"); d["integrator"] = mis
d["liver"]["interior"] = {"type": "homogeneous", "sigma_t": 0.8, "albedo": {"type": "rgb", "value": [0.92, 0.45, 0.35]}}
run("mis homogeneous", d)
