"""Small driver for ncu captures: upload one workload scene and render it once or a few times.
usage: python tools/prof_render.py [c3|c3xl|c5|c2|c1] [res] [spp] [tracking] [repeats]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes

w = sys.argv[1] if len(sys.argv) > 1 else "c3"; res = int(sys.argv[2]) if len(sys.argv) > 2 else 1080
spp = int(sys.argv[3]) if len(sys.argv) > 3 else 1; tracking = sys.argv[4] if len(sys.argv) > 4 else "reference"
rep = int(sys.argv[5]) if len(sys.argv) > 5 else 1
if w == "c1": d = scenes.cornell_box_path(res=res, spp=spp)
elif w == "c2": d = scenes.cornell_box_fog(res=res, spp=spp)
elif w == "c5": d = scenes.sphere_liver_inverse(res=res, spp=spp, grid_res=128, tracking=tracking)
else: d = scenes.liver_single_mesh(res=res, spp=spp, grid_res=512 if w == "c3xl" else 256, tracking=tracking,
                                   surface=os.environ.get("LR_SURFACE", "roughdielectric"))   # LR_SURFACE=bumpmap: the fork's own boundary
if os.environ.get("LR_CELL"): d["integrator"]["majorant_cell"] = int(os.environ["LR_CELL"])
sc = mi.load_dict(d); ctx = mi.Context(0).upload(sc)
for r in range(rep):
    t = time.time()
    if w == "c5":
        img = ctx.render(seed=r, spp=spp); g = np.ones_like(img) / img.size
        ctx.render_backward(g, seed=r + 1, spp=spp)
    else:
        img = ctx.render(seed=r, spp=spp)
    st = ctx.stats(); dt = st["device_ms"]
    print(f"{w} res {res} spp {spp} {tracking}: {dt:.2f} ms  {st['samples'] / dt / 1e3:.2f} Msamples/s  {st['collisions'] / dt / 1e3:.1f} Mcoll/s "
          f"{st['rays'] / dt / 1e3:.1f} Mrays/s  coll/sample {st['collisions'] / max(st['samples'], 1):.2f}  track+nee {st['tracking_ms']:.1f} ms/{st['tracking_launches']}  mean {float(img[..., :3].mean()):.5f}", flush=True)
