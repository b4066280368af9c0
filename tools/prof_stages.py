"""Per-stage device time of the forward wavefront (lr_profile_enable: one batch in flight, CUDA events around each stage).
usage: python tools/prof_stages.py [workload c3|c3xl|c2|c4] [res] [spp] [tracking]"""
import os, sys, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes

w = sys.argv[1] if len(sys.argv) > 1 else "c3"
res = int(sys.argv[2]) if len(sys.argv) > 2 else 1080
spp = int(sys.argv[3]) if len(sys.argv) > 3 else 64
tracking = sys.argv[4] if len(sys.argv) > 4 else "dda"
if w == "c2":
    d = scenes.cornell_box_fog(res=res, spp=spp)
elif w == "c4":
    d = scenes.liver_multi_mesh(res=res, spp=spp, tracking=tracking)
else:
    d = scenes.liver_single_mesh(res=res, spp=spp, grid_res=512 if w == "c3xl" else 256, tracking=tracking)
sc = mi.load_dict(d); ctx = mi.Context(0).upload(sc)
ctx.film_clear(); ctx.render_lanes(1, spp, 0, 0)                     # warm-up
ctx.film_clear(); ctx.render_lanes(2, spp, 0, 0); st0 = ctx.stats()
ctx.profile(True); ctx.film_clear(); ctx.render_lanes(2, spp, 0, 0); st = ctx.stats(); prof = ctx.stage_profile(); ctx.profile(False)
tot = sum(v["ms"] for v in prof.values())
print(f"{w} {res}x{res} {spp} spp {tracking}: overlapped {st0['device_ms']:.2f} ms = {st0['samples'] / st0['device_ms'] / 1e3:.1f} Msamples/s; "
      f"profile mode (1 batch in flight) {st['device_ms']:.2f} ms, sum of stages {tot:.2f} ms")
print(f"rays/sample {st['rays'] / st['samples']:.3f} collisions/sample {st['collisions'] / st['samples']:.3f} launches {st['launches']}")
for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms"]):
    if v["launches"]:
        print(f"  {k:18s} {v['ms']:9.3f} ms {100 * v['ms'] / tot:5.1f} %  launches {v['launches']:5d}  avg {1e3 * v['ms'] / v['launches']:8.1f} us")
