import sys, time
sys.path.insert(0, "/root/repo")
import numpy as np
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes
for res, spp in ((256, 16), (1080, 1), (1080, 16), (1080, 64)):
    sc = mi.load_dict(scenes.liver_single_mesh_bio(res=res, spp=spp))
    ctx = mi.Context(0).upload(sc); ctx.render(seed=0, spp=spp)
    t = time.perf_counter(); ctx.render(seed=1, spp=spp); dt = time.perf_counter() - t
    print(res, spp, round(dt * 1e3, 2), "ms", round(res * res * spp / dt / 1e6, 1), "M/s", flush=True)
# volpath per-lane kernel for comparison
import os
os.environ["LR_MEGAKERNEL"] = "1"
sc = mi.load_dict(scenes.liver_single_mesh(res=1080, spp=16, grid_res=128, tracking="reference"))
ctx = mi.Context(0).upload(sc); ctx.render(seed=0, spp=16)
t = time.perf_counter(); ctx.render(seed=1, spp=16); dt = time.perf_counter() - t
print("volpath per-lane", round(dt * 1e3, 2), "ms", round(1080 * 1080 * 16 / dt / 1e6, 1), "M/s")
