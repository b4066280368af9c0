import sys, time, os
sys.path.insert(0, "/root/repo")
import numpy as np
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes
for spp in (64, 8):
    sc = mi.load_dict(scenes.sphere_liver_inverse(res=512, spp=spp, grid_res=128, tracking="dda"))
    ctx = mi.Context(0).upload(sc)
    img = ctx.render(seed=0, spp=spp); g = np.ones_like(img) / img.size
    ctx.render_backward(g, seed=1, spp=spp)
    import torch
    gd = torch.from_numpy(g).cuda()
    for r in range(3):
        torch.cuda.synchronize(); t = time.perf_counter(); ctx.render_lanes(r, spp, 0, 0); torch.cuda.synchronize(); t1 = time.perf_counter()
        ctx.render_backward(gd, seed=r + 1, spp=spp); torch.cuda.synchronize(); t2 = time.perf_counter()
        st = ctx.stats()
        print(f"spp {spp}: primal {1e3*(t1-t):.2f} ms backward {1e3*(t2-t1):.2f} ms launches {st['launches']} device_ms {st['device_ms']:.2f}", flush=True)
    ctx.profile(True); ctx.render_backward(gd, seed=9, spp=spp); p = ctx.stage_profile(); ctx.profile(False)
    print({k: (round(v["ms"], 2), v["launches"]) for k, v in p.items() if v["launches"]})
