import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes, multi
sc = mi.load_dict(scenes.sphere_liver_inverse(res=512, spp=64, grid_res=128, tracking="dda")); ctx = mi.Context(0).upload(sc)
if len(sys.argv) > 1 and sys.argv[1] == "torchstream": ctx.set_stream(torch.cuda.current_stream().cuda_stream)
img = torch.empty((512, 512, 3), device="cuda"); target = torch.full_like(img, 0.5)
for k in range(8):
    t0 = time.perf_counter(); ctx.film_clear(); ctx.render_lanes(k, 64, 0, 0); s1 = ctx.stats(); t1 = time.perf_counter()
    ctx.film_develop(img); gg = (2.0 / img.numel()) * (img - target); torch.cuda.synchronize(); t2 = time.perf_counter()
    ctx.grid_grad_zero(0); ctx.render_backward(gg, seed=int(mi.sample_tea_32(k, 1)[0]), spp=64); s2 = ctx.stats(); t3 = time.perf_counter()
    print(f"step {k}: primal dev {s1['device_ms']:.2f} wall {1e3 * (t1 - t0):.2f} launches {s1['launches']} | backward dev {s2['device_ms']:.2f} wall {1e3 * (t3 - t2):.2f} launches {s2['launches']}", flush=True)
