"""Time the fork's own scenes (biovolpath + tissue media) on the GPU and, bounded, on the CPU oracle."""
import sys, time, json, os
sys.path.insert(0, "/root/repo")
import numpy as np
import liverrenderer_b200 as mi
from liverrenderer_b200 import scenes

res = int(sys.argv[1]) if len(sys.argv) > 1 else 1080
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 256
out = {}
for name, b, smp in (("single", scenes.liver_single_mesh_bio, "independent"), ("multi", scenes.liver_multi_mesh_bio, "ldsampler")):
    sc = mi.load_dict(b(res=res, spp=spp, sampler=smp))
    ctx = mi.Context(0).upload(sc)
    ctx.render(seed=0, spp=4)
    ts = []
    for i in range(3):
        t = time.perf_counter(); img = ctx.render(seed=i + 1, spp=spp); ts.append(time.perf_counter() - t)
    n = res * res * spp
    st = ctx.stats()
    out[name] = {"ms": [round(x * 1e3, 2) for x in ts], "Msamples_s": round(n / min(ts) / 1e6, 1), "mean": img[..., :3].mean((0, 1)).tolist(),
                 "rays_per_sample": st["rays"] / max(st["samples"], 1), "finite": bool(np.isfinite(img).all())}
    print(name, out[name], flush=True)
if "--cpu" in sys.argv:
    from tests import orc
    sc = mi.load_dict(scenes.liver_single_mesh_bio(res=res, spp=spp))
    o = orc.OracleScene(sc, fast=True); n = 2_000_000
    t = time.perf_counter(); o.sample_lanes(0, spp, 0, n); dt = time.perf_counter() - t
    out["cpu_single"] = {"Msamples_s": round(n / dt / 1e6, 3), "threads": os.cpu_count()}
    print(out["cpu_single"])
os.makedirs("/root/repo/gpurun_out", exist_ok=True)
json.dump(out, open("/root/repo/gpurun_out/bench_bio.json", "w"), indent=1)
