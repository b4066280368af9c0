"""Time volpathmis (per-lane kernel) on the C3 frame, both MIS modes, next to the volpath wavefront."""
import sys, time, json, os
sys.path.insert(0, "/root/repo")
import numpy as np
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes
res, spp = 1080, 16
out = {}
for name, integ in (("volpathmis_spectral", {"type": "volpathmis", "max_depth": 12, "rr_depth": 5}),
                    ("volpathmis_plain", {"type": "volpathmis", "max_depth": 12, "rr_depth": 5, "use_spectral_mis": False}),
                    ("volpath_wavefront_reference_order", None)):
    d = scenes.liver_single_mesh(res=res, spp=spp, grid_res=256, tracking="reference")
    if integ: d["integrator"] = integ
    ctx = mi.Context(0).upload(mi.load_dict(d)); ctx.render(seed=0, spp=1)
    t = time.perf_counter(); img = ctx.render(seed=1, spp=spp); dt = time.perf_counter() - t
    st = ctx.stats()
    out[name] = {"ms": round(dt * 1e3, 1), "Msamples_s": round(res * res * spp / dt / 1e6, 2), "collisions_per_sample": st["collisions"] / st["samples"],
                 "mean": float(img[..., :3].mean()), "finite": bool(np.isfinite(img).all())}
    print(name, out[name], flush=True)
os.makedirs("/root/repo/gpurun_out", exist_ok=True)
json.dump(out, open("/root/repo/gpurun_out/bench_mis.json", "w"), indent=1)
