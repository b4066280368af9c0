import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes, multi
sc = mi.load_dict(scenes.sphere_liver_inverse(res=512, spp=64, grid_res=128, tracking="dda")); ctx = mi.Context(0).upload(sc)
g = torch.full((512, 512, 3), 1.0 / (512 * 512 * 3), device="cuda")
for k in (3, 2, 5):
    seed = int(mi.sample_tea_32(k, 1)[0])
    ctx.profile(True); ctx.grid_grad_zero(0); ctx.render_backward(g, seed=seed, spp=64); p = ctx.stage_profile(); st = ctx.stats()
    print(k, seed, f"{st['device_ms']:.1f} ms", {n: (round(v["ms"], 2), v["launches"]) for n, v in p.items() if v["ms"] > 0.3}, flush=True)
ctx.profile(False)
for k in (0, 1, 2):
    ctx.profile(True); ctx.film_clear(); ctx.render_lanes(k, 64, 0, 0); p = ctx.stage_profile(); st = ctx.stats()
    print("primal", k, f"{st['device_ms']:.1f} ms", {n: (round(v["ms"], 2), v["launches"]) for n, v in p.items() if v["ms"] > 0.3}, flush=True)
