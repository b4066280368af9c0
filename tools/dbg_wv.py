"""Bring-up check of the forward wavefront: small C3-like scene, wavefront vs per-lane kernel, both tracking modes.
usage: python tools/dbg_wv.py [res] [spp] [grid]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes

res = int(sys.argv[1]) if len(sys.argv) > 1 else 32
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 4
grid = int(sys.argv[3]) if len(sys.argv) > 3 else 16
n = res * res * spp
for tracking in ("reference", "dda"):
    sc = mi.load_dict(scenes.liver_single_mesh(res=res, spp=spp, grid_res=grid, tracking=tracking))
    os.environ.pop("LR_MEGAKERNEL", None)
    wf = mi.Context(0).upload(sc); a = wf.sample_lanes(3, spp, 0, n); sa = wf.stats()
    os.environ["LR_MEGAKERNEL"] = "1"
    mk = mi.Context(0).upload(sc); b = mk.sample_lanes(3, spp, 0, n); sb = mk.stats()
    os.environ.pop("LR_MEGAKERNEL", None)
    err = np.abs(a[:, :3] - b[:, :3]) / np.maximum(np.abs(b[:, :3]), 1e-3)
    print(tracking, "match 1e-5:", float((err.max(axis=1) < 1e-5).mean()), "valid eq:", bool(np.array_equal(a[:, 3], b[:, 3])),
          "mean", a[:, :3].mean(), b[:, :3].mean())
    print("  wf", {k: sa[k] for k in ("samples", "rays", "collisions", "real_scatters", "launches")})
    print("  mk", {k: sb[k] for k in ("samples", "rays", "collisions", "real_scatters", "launches")})
    img = wf.render(seed=3, spp=spp)
    print("  film mean", img[..., :3].mean(), "lanes mean", a[:, :3].mean(), "alpha", img[..., 3].min(), img[..., 3].max())
