"""Driver for ncu captures of the per-lane integrators of SURVEY.md 8(f): one render of
  single | multi : the fork's Liver-SingleMesh / Liver-MultiMesh scene under biovolpath
  mis            : the C3 frame under volpathmis (spectral MIS), reference-order tracking
usage: python tools/prof_bio.py [single|multi|mis] [res] [spp]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes
w = sys.argv[1] if len(sys.argv) > 1 else "single"; res = int(sys.argv[2]) if len(sys.argv) > 2 else 512; spp = int(sys.argv[3]) if len(sys.argv) > 3 else 64
if w == "mis":
    d = scenes.liver_single_mesh(res=res, spp=spp, grid_res=256, tracking="reference"); d["integrator"] = {"type": "volpathmis", "max_depth": 12, "rr_depth": 5}
else:
    d = (scenes.liver_single_mesh_bio if w == "single" else scenes.liver_multi_mesh_bio)(res=res, spp=spp)
ctx = mi.Context(0).upload(mi.load_dict(d))
img = ctx.render(seed=1, spp=spp); st = ctx.stats(); dt = st["device_ms"]
print(f"{w} res {res} spp {spp}: {dt:.2f} ms  {st['samples'] / dt / 1e3:.2f} Msamples/s  {st['rays'] / dt / 1e3:.1f} Mrays/s  rays/sample {st['rays'] / st['samples']:.2f}  "
      f"coll/sample {st['collisions'] / st['samples']:.2f}  mean {float(img[..., :3].mean()):.5f}", flush=True)
