import sys, time, os
sys.path.insert(0, "/root/repo")
import numpy as np
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes
sc = mi.load_dict(scenes.sphere_liver_inverse(res=512, spp=64, grid_res=128, tracking="dda"))
ctx = mi.Context(0).upload(sc)
img = ctx.render(seed=0, spp=64); g = np.ones_like(img) / img.size
ctx.render_backward(g, seed=1, spp=64)
for r in range(3):
    ctx.grid_grad_zero(0)
    t = time.perf_counter(); img = ctx.render(seed=r, spp=64); t1 = time.perf_counter(); ctx.render_backward(g, seed=r + 1, spp=64); t2 = time.perf_counter()
    print("fuse_off" if os.environ.get("LR_TRACK_FUSE_OFF") else "fused", "primal %.2f ms backward %.2f ms step %.2f ms" % ((t1 - t) * 1e3, (t2 - t1) * 1e3, (t2 - t) * 1e3), flush=True)
