"""Diagnostic: grid gradient behind a bump-mapped boundary -- wavefront PRB vs per-lane kernel (LR_MEGAKERNEL=1) vs oracle."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes
from tests import orc

scale = float(sys.argv[1]) if len(sys.argv) > 1 else 0.05
res, spp = 16, 4
d = scenes.sphere_liver_inverse(res=res, spp=spp, grid_res=8)
d["sphere"]["interior"]["sigma_t"]["max_value"] = 0.5
d["sphere"]["bsdf"] = {"type": "bumpmap", "scale": scale, "tex": {"type": "bitmap", "bitmap": scenes.tissue_texture(64), "raw": True}, "nested": dict(d["sphere"]["bsdf"])}
sc = mi.load_dict(d); g = np.ones((res, res, 3), dtype=np.float32); n = res * res * spp
o = orc.OracleScene(sc); o.render_backward(g, seed=1, spp=spp); gcpu = o.grid_grad_read(0)
cpu_l = o.sample_lanes(0, spp, 0, n)
out = {}
for mk in ("0", "1"):
    os.environ["LR_MEGAKERNEL"] = mk
    ctx = mi.Context(0).upload(sc); ctx.render_backward(g, seed=1, spp=spp); out[mk] = ctx.grid_grad_read(0)
    gl = ctx.sample_lanes(0, spp, 0, n)
    bad = np.abs(gl[:, :3] - cpu_l[:, :3]).max(axis=1) / np.maximum(np.abs(cpu_l[:, :3]).max(axis=1), 1e-3) > 1e-3
    print(f"scale {scale} LR_MEGAKERNEL={mk}: |g_gpu - g_cpu| / |g_cpu| = {np.linalg.norm(out[mk] - gcpu) / np.linalg.norm(gcpu):.3e}; primal lanes off: {int(bad.sum())} of {n}")
    ctx.close()
print(f"wavefront vs per-lane: {np.linalg.norm(out['0'] - out['1']) / np.linalg.norm(out['1']):.3e};  |g_cpu| = {np.linalg.norm(gcpu):.3f}, max |g_cpu| = {np.abs(gcpu).max():.3f}")
