import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes
sc = mi.load_dict(scenes.sphere_liver_inverse(res=512, spp=64, grid_res=128, tracking="dda")); ctx = mi.Context(0).upload(sc)
g = torch.full((512, 512, 3), 1.0 / (512 * 512 * 3), device="cuda")
img = torch.empty((512, 512, 3), device="cuda"); ctx.render(seed=0, spp=64, out=img)
gg = (2.0 / img.numel()) * (img - 0.5)
print("grad_in stats: const", float(g[0, 0, 0]), " bench: min", float(gg.min()), "max", float(gg.max()), "zeros", int((gg == 0).sum()))
for rep in range(3):
    for name, gi in (("const", g), ("bench", gg), ("abs(bench)", gg.abs()), ("bench*1e3", gg * 1e3)):
        ctx.grid_grad_zero(0); ctx.render_backward(gi, seed=3, spp=64); print(rep, name, f"{ctx.stats()['device_ms']:.2f} ms", flush=True)
ctx.profile(True)
for name, gi in (("const", g), ("bench", gg)):
    ctx.profile(True); ctx.grid_grad_zero(0); ctx.render_backward(gi, seed=3, spp=64); p = ctx.stage_profile()
    print(name, {k: round(v["ms"], 2) for k, v in p.items() if v["ms"] > 0.05})
