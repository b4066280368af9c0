"""C5 backward timing under the conditions bench.py runs it: plain, with the (identity) sample-share lane map, on torch's stream, after a hot C3 load."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes
sc = mi.load_dict(scenes.sphere_liver_inverse(res=512, spp=64, grid_res=128, tracking="dda")); ctx = mi.Context(0).upload(sc)
g = torch.full((512, 512, 3), 1.0 / (512 * 512 * 3), device="cuda")
def t(label, n=3):
    for k in range(n):
        ctx.grid_grad_zero(0); ctx.render_backward(g, seed=k + 1, spp=64); ms = ctx.stats()["device_ms"]
    print(f"{label}: {ms:.2f} ms", flush=True)
t("plain")
ctx.set_lane_map(64, 0, 64); t("lane map (64, 0, 64)"); ctx.set_lane_map()
ctx.set_stream(torch.cuda.current_stream().cuda_stream); t("torch stream")
c3 = mi.Context(0).upload(mi.load_dict(scenes.liver_single_mesh(res=1080, spp=1024, grid_res=256, tracking="dda")))
c3.film_clear(); c3.render_lanes(0, 1024, 0, 0); c3.film_clear(); c3.render_lanes(1, 1024, 0, 0)
t("after 2.3 s of C3 (pools of both contexts resident)")
os.system("nvidia-smi --query-gpu=clocks.sm,power.draw,memory.used --format=csv,noheader")
# the bench's own step
from liverrenderer_b200 import multi
img = torch.empty((512, 512, 3), device="cuda"); target = torch.full_like(img, 0.5)
for k in range(3):
    multi.render_distributed(ctx, k, 64, 0, 1, mode="samples", film_buf=None, out=img); ms_p = ctx.stats()["device_ms"]
    gg = (2.0 / img.numel()) * (img - target)
    multi.backward_distributed(ctx, gg, mi.sample_tea_32(k, 1)[0], 64, 0, 1, grad_buf=True, mode="samples"); st = ctx.stats()
    print(f"bench-like step {k}: primal {ms_p:.2f} ms, backward {st['device_ms']:.2f} ms, collisions {st['collisions']}, rays {st['rays']}", flush=True)
ctx.grid_grad_zero(0); ctx.render_backward(gg, seed=5, spp=64); print("backward with the bench's grad_in, small seed:", ctx.stats()["device_ms"])
ctx.grid_grad_zero(0); ctx.render_backward(g, seed=int(mi.sample_tea_32(1, 1)[0]), spp=64); print("backward with constant grad_in, TEA seed:", ctx.stats()["device_ms"], ctx.stats()["collisions"])
