import os, sys, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes
def scene(albedo, max_depth, spp=64, res=24):
    rng = np.random.default_rng(12345); dens = (0.02 + 0.05 * rng.random((8, 8, 8, 1))).astype(np.float32)
    d = scenes.sphere_liver_inverse(res=res, spp=spp, grid_res=8, density=dens)
    d["sphere"]["interior"]["sigma_t"]["max_value"] = 0.08
    d["sphere"]["interior"]["albedo"]["value"] = [albedo] * 3
    d["integrator"]["max_depth"] = max_depth
    return mi.load_dict(d)
rng = np.random.default_rng(3); g = rng.standard_normal((24, 24, 3)).astype(np.float32)
for albedo, md in [(0.0, 8), (0.8, 2), (0.8, 3), (0.8, 8)]:
    sc = scene(albedo, md); out = {}
    for mk in ("0", "1"):
        os.environ["LR_MEGAKERNEL"] = mk
        ctx = mi.Context(0).upload(sc); ctx.grid_grad_zero(0); ctx.render_backward(g, seed=7, spp=64); out[mk] = ctx.grid_grad_read(0).astype(np.float64)
    a, b = out["0"], out["1"]
    i = np.argsort(-np.abs(a - b))[:4]
    print(f"albedo {albedo} max_depth {md}: rel L2 {np.linalg.norm(a - b) / np.linalg.norm(b):.4f} sum wf {a.sum():.5f} mk {b.sum():.5f}  |b| {np.linalg.norm(b):.4f} top diffs", [(int(k), round(a[k], 4), round(b[k], 4)) for k in i])
