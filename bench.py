#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 volumetric path-tracing backend (BASELINE.json metric:
samples/s + Mrays/s of the `volpath` forward render and the `prbvolpath` gradient step).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--partition samples|seeds]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one full render of the workload scene (C3: Liver-SingleMesh, heterogeneous gridvolume, `volpath`,
1080x1080, --spp samples per pixel) with every input already resident in HBM.
Multi-GPU, one process per GPU.  `--partition samples` (default, "scaling": "strong"): ONE frame is split by sample index --
rank r renders spp / N samples of every pixel (lr_set_lane_map), no communication while rendering, then the film accumulators are
summed IN PLACE (NCCL all-reduce on lr_film_device_ptr) and developed; `value` = samples of the frame / max-over-ranks time.
`--partition seeds` ("weak"): every rank renders the whole config with seed + rank; at N > 1 the default run reports it under
the key `weak`.  The `prb` object is the C5 inverse step split the same way (primal + prbvolpath adjoint + in-place NCCL
all-reduce of the grid gradient).  `e2e` is the same render through the public host API (`params.update()` + render) with pinned
HOST buffers: the density grid goes host->device and the developed image device->host inside the timed region.
`stages` is the per-stage table of one extra step in the library's profile mode (one batch in flight, CUDA events around every
launch on the launching stream); `rooflines` bounds each dominant stage, `roofline` is the entry of the largest one.
`--impl reference` times the CPU oracle (the restatement of the reference's algorithm, oracle/, rebuilt -O3 -march=native on the
box) on the host cores.  Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

# stdout carries exactly one JSON line: file descriptor 1 is pointed at stderr for the whole run (NCCL prints its banner /
# debug lines straight to fd 1 when the box sets NCCL_DEBUG) and the line is written to the saved descriptor at the end.
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(line: dict) -> None:
    sys.stdout.flush()
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "volpath_fwd_samples_per_s"
UNIT = "samples/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", choices=["c1", "c2", "c3", "c3xl"])
    ap.add_argument("--res", type=int, default=1080)
    ap.add_argument("--spp", type=int, default=0, help="samples per pixel per step (0 => the config's)")
    ap.add_argument("--tracking", default="dda", choices=["reference", "dda"])
    ap.add_argument("--partition", default="samples", choices=["samples", "seeds"],
                    help="N > 1: samples = one frame split by sample index (strong scaling), seeds = every rank renders the frame (weak)")
    ap.add_argument("--no-extras", action="store_true", help="skip the reference-order / bump-map / C3-XL side measurements")
    ap.add_argument("--no-prb", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-spp", type=int, default=4, help="spp of the bounded CPU-baseline sample")
    return ap.parse_args()


def workload_scene(args, for_cpu=False):
    from liverrenderer_b200 import scenes
    w = args.workload; res = args.res
    if w == "c1":
        spp = args.spp or 64
        return scenes.cornell_box_path(res=256 if res == 1080 else res, spp=spp), spp, "C1 mi.cornell_box() 256x256 path"
    if w == "c2":
        spp = args.spp or 256
        return scenes.cornell_box_fog(res=res, spp=spp), spp, f"C2 cornell-box fog homogeneous volpath {res}x{res}"
    grid = 512 if w == "c3xl" else 256
    spp = args.spp or 1024
    tracking = "reference" if for_cpu else args.tracking
    d = scenes.liver_single_mesh(res=res, spp=spp, grid_res=grid, tracking=tracking)
    return d, spp, f"C3 Liver-SingleMesh heterogeneous gridvolume {grid}^3 volpath {res}x{res}"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows = []; self.p = None; self.index = index

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True); self.t.start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=2)
        except Exception:
            self.p.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); pw.append(float(r[2]))
            except ValueError:
                continue
            for k, n in enumerate(names):
                if r[3 + k].lower().startswith("active"):
                    reasons.add(n)
        # under load = the upper half of the power samples
        if sm:
            thr = np.median(pw); load = [s for s, p in zip(sm, pw) if p >= thr] or sm
            return {"sm_mhz": float(np.median(load)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                    "power_w_max": float(max(pw)), "samples": len(sm)}
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}


# --------------------------------------------------------------------------------------------------
# CPU side: the oracle (restatement of the reference's llvm_ad_rgb algorithm; the reference itself
# cannot be built or imported here, DESIGN.md "Reference arm") on the box's host cores
# --------------------------------------------------------------------------------------------------
def cpu_sample(args, spp_cpu: int, repeats: int = 1):
    """Renders the workload scene at `spp_cpu` spp with the oracle, all host threads.  Returns (samples/s, dict)."""
    import liverrenderer_b200 as mi
    from tests import orc
    d, _, name = workload_scene(args, for_cpu=True)
    d["sensor"]["sampler"]["sample_count"] = spp_cpu
    sc = mi.load_dict(d); o = orc.OracleScene(sc, threads=0, fast=True)
    cores = os.cpu_count() or 1
    best = None
    for r in range(repeats):
        o.stats_reset(); t = time.perf_counter(); o.render(seed=r, spp=spp_cpu); dt = time.perf_counter() - t
        st = o.stats()
        if best is None or dt < best[0]:
            best = (dt, st)
    dt, st = best
    W, H = sc.film_size
    return st["samples"] / dt, {"value": st["samples"] / dt, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"{name} at {spp_cpu} spp ({W * H * spp_cpu} samples, {dt:.1f} s), oracle/ rebuilt on this host with g++ {orc.FAST_FLAGS}, "
                                          f"std::thread x{cores}, global-majorant delta tracking as the reference",
                                "mrays_per_s": st["rays"] / dt / 1e6, "collisions_per_s": st["collisions"] / dt}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    d, spp, name = workload_scene(args, for_cpu=True)
    import liverrenderer_b200 as mi
    from tests import orc
    spp_cpu = args.cpu_spp
    d["sensor"]["sampler"]["sample_count"] = spp_cpu
    sc = mi.load_dict(d); o = orc.OracleScene(sc, threads=0, fast=True); W, H = sc.film_size
    cores = os.cpu_count() or 1
    for w in range(min(args.warmup, 1)):                      # one warm-up pass is enough for a CPU code (page-in, thread pool)
        o.render(seed=1000 + w, spp=spp_cpu)
    o.stats_reset(); t = time.perf_counter()
    for k in range(args.steps):
        o.render(seed=k, spp=spp_cpu)
    dt = time.perf_counter() - t; st = o.stats()
    v = st["samples"] / dt
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "strong" if args.partition == "samples" else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": {"workload": name, "spp_per_step": spp_cpu, "samples_per_step": W * H * spp_cpu,
                                            "note": "each step is a bounded sample (reduced spp) of the same scene"},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"{args.steps} x ({W}x{H} at {spp_cpu} spp) through the oracle rebuilt on this host (g++ {orc.FAST_FLAGS}) on {cores} threads"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "mrays_per_s": st["rays"] / dt / 1e6, "collisions_per_s": st["collisions"] / dt, "gpu_launches": 0}
    emit(line)


def cpu_sample_prb(args, res: int = 512, spp: int = 1):
    """C5 on the host cores: primal + prbvolpath adjoint of the oracle (timing build), a bounded sample of the step."""
    import liverrenderer_b200 as mi
    from liverrenderer_b200 import scenes
    from tests import orc
    sc = mi.load_dict(scenes.sphere_liver_inverse(res=res, spp=spp, grid_res=128, tracking="reference"))
    o = orc.OracleScene(sc, threads=0, fast=True); cores = os.cpu_count() or 1
    g = np.full((res, res, 3), 2.0 / (res * res * 3), dtype=np.float32)
    o.grid_grad_zero(0); o.stats_reset(); t = time.perf_counter()
    o.render(seed=0, spp=spp); o.render_backward(g, seed=1, spp=spp); dt = time.perf_counter() - t
    n = res * res * spp
    return {"value": n / dt, "unit": "grad-samples/s", "cores": cores, "kind": "port",
            "sample": f"C5 at {res}x{res} x {spp} spp: oracle primal render + render_backward ({n} lanes, {dt:.1f} s), g++ {orc.FAST_FLAGS}, std::thread x{cores}"}


# --------------------------------------------------------------------------------------------------
# B200 arm
# --------------------------------------------------------------------------------------------------
TRACE_BYTES_PER_RAY = 48          # 32 B in (origin + maxt, direction) + 16 B out (t, prim, u, v): lr_trace.cuh
TRACK_BYTES_PER_COLLISION = 32    # the 8 fp32 corners of one trilinear sigma_t tap
PRB_BYTES_PER_COLLISION = 64      # the same tap read + its 8 gradient atomics


def _load_json(name):
    try:
        return json.load(open(os.path.join(ROOT, "profiles", name)))
    except Exception:
        return None


def timed_render(torch, ctx, stream, barrier, seed, spp, reps=1):
    """One warm render, then `reps` timed ones (CUDA events on the context's stream); returns (ms per render, stats of the last)."""
    ctx.film_clear(); ctx.render_lanes(seed, spp, 0, 0); barrier()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True); e0.record(stream)
    for k in range(reps):
        ctx.film_clear(); ctx.render_lanes(seed + 1 + k, spp, 0, 0)
    e1.record(stream); barrier()
    return e0.elapsed_time(e1) / reps, ctx.stats()


def stage_table(ctx, seed, spp, hbm_peak, ncu):
    """One extra step in profile mode -> per-stage ms / launches / items, and the rooflines built on it."""
    ctx.profile(True); ctx.film_clear(); ctx.render_lanes(seed, spp, 0, 0); st = ctx.stats(); prof = ctx.stage_profile(); ctx.profile(False)
    total = sum(v["ms"] for v in prof.values()) or 1.0
    stages = {k: {"ms": round(v["ms"], 3), "share": round(v["ms"] / total, 4), "launches": v["launches"], "items": v["items"]}
              for k, v in prof.items() if v["launches"] or v["items"]}
    def ncu_of(kernel):
        return (ncu or {}).get("kernels", {}).get(kernel)
    def issue_entry(name, kernel, what):
        s = prof.get(name)
        if not s or not s["ms"]:
            return None
        per_s = s["items"] / (s["ms"] * 1e-3); gbs = per_s * TRACE_BYTES_PER_RAY / 1e9; m = ncu_of(kernel) or {}
        e = {"kernel": kernel, "stage": name, "what": what, "bound": "sm_issue", "share_of_step": round(s["ms"] / total, 4),
             "achieved": per_s / 1e6, "unit": "Mrays/s", "kernel_ms": s["ms"] / max(s["launches"], 1), "units_per_launch": s["items"] / max(s["launches"], 1),
             "algorithmic_bytes_per_unit": TRACE_BYTES_PER_RAY, "algorithmic_GBps": gbs, "frac_of_hbm_peak": gbs / hbm_peak,
             "issue_active_pct": m.get("issue_active_pct"), "threads_per_inst": m.get("threads_per_inst"),
             "lane_issue_utilisation": (m["issue_active_pct"] / 100.0 * m["threads_per_inst"] / 32.0) if m.get("issue_active_pct") and m.get("threads_per_inst") else None,
             "traffic": m.get("dram_bytes_per_launch"), "source": (ncu or {}).get("source")}
        return e
    rl = []
    for name, kernel, what in (("trace_closest", "k_wv_trace<0>", "closest-hit query of every path segment, 4-wide quantized BVH in shared memory"),
                               ("trace_shadow", "k_wv_trace<1>", "any-hit shadow rays of emitter sampling")):
        e = issue_entry(name, kernel, what)
        if e:
            rl.append(e)
    tn_ms = prof.get("track", {}).get("ms", 0.0) + prof.get("nee", {}).get("ms", 0.0)
    if tn_ms > 0:
        b = st["collisions"] * float(TRACK_BYTES_PER_COLLISION); ach = b / (tn_ms * 1e-3) / 1e9
        n_l = prof["track"]["launches"] + prof["nee"]["launches"]; m = ncu_of("k_wv_track_dda") or {}
        rl.append({"kernel": "k_wv_track_dda + k_wv_nee", "stage": "track+nee", "what": "majorant-DDA free-flight / ratio tracking, trilinear sigma_t taps", "bound": "hbm",
                   "share_of_step": round(tn_ms / total, 4), "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                   "algorithmic_bytes_per_unit": TRACK_BYTES_PER_COLLISION, "units_per_launch": st["collisions"] / max(n_l, 1), "kernel_ms": tn_ms / max(n_l, 1),
                   "traffic": m.get("dram_bytes_per_launch"), "l2_GBps": m.get("l2_GBps"), "source": (ncu or {}).get("source")})
    if ncu and ncu.get("dram_bytes_per_sample"):
        ach = ncu["dram_bytes_per_sample"] * st["samples"] / (total * 1e-3) / 1e9
        rl.append({"kernel": "all stages", "stage": "step", "what": "path state + queues + grid + film: DRAM bytes of every kernel of the step (ncu) / step time", "bound": "hbm",
                   "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak, "dram_bytes_per_sample": ncu["dram_bytes_per_sample"], "source": ncu.get("source")})
    return {"stages": stages, "total_ms": total, "samples": st["samples"], "rays": st["rays"], "collisions": st["collisions"]}, rl


def run_b200(args):
    import torch
    import torch.distributed as dist
    import liverrenderer_b200 as mi
    from liverrenderer_b200 import multi

    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py: no CUDA device -- the B200 backend has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local); dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0)); peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650 GB/s (B200_PROFILING.md)"

    d, spp, name = workload_scene(args)
    sc = mi.load_dict(d); W, H = sc.film_size
    ctx = mi.Context(local).upload(sc)
    stream = torch.cuda.current_stream(dev); ctx.set_stream(stream.cuda_stream)
    film = multi.film_tensor(ctx, dev)                                      # aliases lr_film_device_ptr: the all-reduce runs in place
    out_dev = torch.empty((H, W, sc.out_channels), dtype=torch.float32, device=dev)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)          # > 126 MB L2
    _, spp_pass, _ = multi.lane_count(W, H, spp)

    def set_partition(mode):
        share = multi.sample_share(spp_pass, rank, world) if (mode == "samples" and world > 1) else None
        if share is not None:
            ctx.set_lane_map(*share)
        else:
            ctx.set_lane_map()
        return share

    def fwd_step(k, mode):
        ctx.film_clear(); ctx.render_lanes(1000 * k + (rank if mode == "seeds" else 0), spp, 0, 0)
        st = ctx.stats()
        if world > 1:
            dist.all_reduce(film)
        ctx.film_develop(out_dev)
        return st

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(mode, steps, warmup, clocks=None):
        set_partition(mode)
        for w in range(warmup):
            fwd_step(-1 - w, mode)
        barrier()
        if clocks:
            clocks.start()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        tot = {"samples": 0, "rays": 0, "collisions": 0, "real_scatters": 0, "launches": 0}
        for k in range(steps):
            flush.fill_(k & 0xff)                                             # L2 flush between timed iterations
            barrier()
            ev[k][0].record(stream); st = fwd_step(k, mode); ev[k][1].record(stream)
            for key in tot:
                tot[key] += st[key]
            tot["launches"] += 1                                               # develop (the all-reduce is NCCL's kernel, not counted)
        barrier(); clk = clocks.stop() if clocks else None
        ms = sum(a.elapsed_time(b) for a, b in ev)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        cnt = torch.tensor([tot["samples"], tot["rays"], tot["collisions"], tot["launches"]], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX); dist.all_reduce(cnt)
        set_partition("seeds")
        samples, rays, colls, launches = (float(x) for x in cnt.tolist())
        return dict(ms=float(t.item()), samples=samples, rays=rays, colls=colls, launches=launches, clk=clk, steps=steps)

    main_run = timed(args.partition, args.steps, args.warmup, ClockSampler(local))
    weak = None
    if world > 1 and args.partition == "samples":
        r = timed("seeds", min(args.steps, 2), 1)
        weak = {"value": r["samples"] / (r["ms"] * 1e-3), "unit": UNIT, "ms_per_step": r["ms"] / r["steps"], "steps": r["steps"],
                "partition": "every rank renders the whole config with seed + rank; film all-reduce in place"}

    # ---- per-stage table + rooflines: one extra step of this rank's share in profile mode ----
    ncu = _load_json("r2_kernel_metrics.json") if (args.workload == "c3" and args.tracking == "dda") else None
    set_partition(args.partition)
    table, rooflines = stage_table(ctx, 4242, spp, hbm_peak, ncu)
    set_partition("seeds"); barrier()
    for e in rooflines:
        e["peak_source"] = peak_src
    dom = max((e for e in rooflines if e["stage"] != "step"), key=lambda e: e.get("share_of_step", 0.0), default=None)
    roofline = None
    if dom is not None:
        gbs = dom.get("algorithmic_GBps", dom.get("achieved")) if dom["bound"] == "sm_issue" else dom["achieved"]
        roofline = {"bound": "hbm", "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak, "traffic": dom.get("traffic"),
                    "kernel": dom["kernel"], "stage": dom["stage"], "share_of_step": dom.get("share_of_step"), "kernel_ms": dom["kernel_ms"],
                    "algorithmic_bytes_per_unit": dom["algorithmic_bytes_per_unit"], "units_per_launch": dom["units_per_launch"],
                    "algorithmic_bytes_per_launch": dom["algorithmic_bytes_per_unit"] * dom["units_per_launch"], "peak_source": peak_src,
                    "limiter": dom["bound"], "lane_issue_utilisation": dom.get("lane_issue_utilisation"), "issue_active_pct": dom.get("issue_active_pct"),
                    "threads_per_inst": dom.get("threads_per_inst"), "traffic_source": dom.get("source"),
                    "timing": "CUDA events around every launch of the stage on its stream, one extra step in profile mode (one batch in flight; "
                              "in the timed region batches overlap on 3 streams, so per-launch intervals there would overlap too)",
                    "note": "the dominant kernel reads its BVH from shared memory: its HBM fraction is small by design and SM issue is the bound "
                            "(north star: 'SM issue rate for traversal'); see `rooflines` for every stage"}

    extras = {}
    if world == 1 and not args.no_extras and args.workload in ("c3", "c3xl") and args.tracking == "dda":
        from liverrenderer_b200 import scenes as _sc
        def side(dd, label, reps=2):
            c2 = mi.Context(local).upload(mi.load_dict(dd)); c2.set_stream(stream.cuda_stream)
            ms, st2 = timed_render(torch, c2, stream, barrier, 7, spp, reps); c2.close()
            return {"value": st2["samples"] / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "mrays_per_s": st2["rays"] / (ms * 1e-3) / 1e6,
                    "collisions_per_sample": st2["collisions"] / max(st2["samples"], 1), "collisions_per_s": st2["collisions"] / (ms * 1e-3), "note": label}
        a2 = argparse.Namespace(**vars(args)); a2.tracking = "reference"
        extras["reference_order_tracking"] = side(workload_scene(a2)[0], "LR_TRACKING_REFERENCE: identical PCG32 draw order to the reference (one global majorant), "
                                                  "the mode of the per-path parity tests and the algorithm the CPU baseline runs", reps=1)
        if args.workload == "c3":
            extras["bumpmap_surface"] = side(_sc.liver_single_mesh(res=args.res, spp=spp, grid_res=256, tracking="dda", surface="bumpmap"),
                                             "C3 with the fork's own boundary BSDF: bumpmap(scale 0.005, bitmap 512^2 rgb) -> dielectric 1.38 (scene.xml:31-40)")
            xl = side(_sc.liver_single_mesh(res=args.res, spp=spp, grid_res=512, tracking="dda"), "C3-XL: the same frame on a 512^3 sigma_t grid (512 MiB fp32 > 126 MB L2)")
            b = xl["collisions_per_s"] * TRACK_BYTES_PER_COLLISION / 1e9
            xl["ratio_to_c3"] = xl["value"] / (main_run["samples"] / (main_run["ms"] * 1e-3))
            xl["roofline"] = {"bound": "hbm", "achieved": b, "peak": hbm_peak, "unit": "GB/s", "frac": b / hbm_peak, "traffic": None,
                              "note": "32 B per collision over the WHOLE step time (tracking is one stage of several), so a lower bound on the tracking kernels' own rate"}
            extras["c3xl"] = xl
            # the fork's own rendering model -- biovolpath + tissue media -- on the two scenes whose wall-clock the
            # reference tree records (unstated hardware and variant, so a ratio to them is context, not a comparison)
            def bio(builder, label, published, cpu_res=540):
                spp_b = 256
                c2 = mi.Context(local).upload(mi.load_dict(builder(res=args.res, spp=spp_b))); c2.set_stream(stream.cuda_stream)
                ms, st2 = timed_render(torch, c2, stream, barrier, 7, spp_b, 2)
                c2.profile(True); c2.film_clear(); c2.render_lanes(11, spp_b, 0, 0); prof = c2.stage_profile(); c2.profile(False); c2.close()
                tot = sum(v["ms"] for v in prof.values()) or 1.0
                stages = {k: {"ms": round(v["ms"], 3), "share": round(v["ms"] / tot, 4), "launches": v["launches"], "items": v["items"],
                              "Mitems_per_s": round(v["items"] / (v["ms"] * 1e-3) / 1e6, 1) if v["ms"] else None} for k, v in prof.items() if v["launches"] or v["items"]}
                tr_ms = sum(prof[k]["ms"] for k in ("trace_closest", "trace_shadow") if k in prof); tr_items = sum(prof[k]["items"] for k in ("trace_closest", "trace_shadow") if k in prof)
                gbs = tr_items * TRACE_BYTES_PER_RAY / (tr_ms * 1e-3) / 1e9 if tr_ms else None
                e = {"stages": stages, "stage_total_ms": round(tot, 3),
                     "roofline": {"bound": "hbm", "limiter": "sm_issue", "kernel": "k_wv_trace<0> + k_wv_trace<2>", "share_of_step": round(tr_ms / tot, 4),
                                  "achieved": gbs, "peak": hbm_peak, "unit": "GB/s", "frac": gbs / hbm_peak if gbs else None, "traffic": None,
                                  "algorithmic_bytes_per_unit": TRACE_BYTES_PER_RAY, "mrays_per_s": tr_items / (tr_ms * 1e-3) / 1e6 if tr_ms else None,
                                  "note": "the TRACE stage is the one the volpath wavefront uses (BVH in shared memory; ncu figures under rooflines[trace_closest]); "
                                          "per-launch times from CUDA events in profile mode (one batch in flight)"},
                     "value": st2["samples"] / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "spp": spp_b, "mrays_per_s": st2["rays"] / (ms * 1e-3) / 1e6,
                     "rays_per_sample": st2["rays"] / max(st2["samples"], 1), "kernel": "wavefront biovolpath (lr_wavefront_bio.cuh: GEN, TRACE, k_bio_track / shade_a / nee / shade_b over queues)", "launches": st2.get("launches"), "note": label,
                     "published_reference": {"value": published, "unit": UNIT, "source": "BASELINE.md (1920x1080, 256 spp; hardware and variant not stated)"}}
                try:
                    from tests import orc
                    o = orc.OracleScene(mi.load_dict(builder(res=cpu_res, spp=64)), threads=0, fast=True)
                    t0 = time.perf_counter(); o.render(seed=0, spp=64); dt = time.perf_counter() - t0
                    e["cpu_baseline"] = {"value": cpu_res * cpu_res * 64 / dt, "unit": UNIT, "cores": os.cpu_count() or 1, "kind": "port",
                                         "sample": f"the same frame at {cpu_res}x{cpu_res}, 64 spp ({cpu_res * cpu_res * 64} samples, {dt:.1f} s), oracle/ -O3 -march=native, std::thread"}
                except Exception as ex:                                   # the oracle is the checker, never a dependency of the measured arm
                    e["cpu_baseline"] = {"unavailable": str(ex)[:120]}
                return e
            # SURVEY.md 8(f) rank 4 (first half): volpathmis on the C3 frame, wavefront, global-majorant tracking in the reference's draw order
            # (it has no production-tracking variant); the comparable volpath figure is extras.reference_order_tracking.
            def mis(spectral):
                dd = _sc.liver_single_mesh(res=args.res, spp=64, grid_res=256, tracking="reference")
                dd["integrator"] = {"type": "volpathmis", "max_depth": 12, "rr_depth": 5, "use_spectral_mis": spectral}
                c2 = mi.Context(local).upload(mi.load_dict(dd)); c2.set_stream(stream.cuda_stream)
                ms, st2 = timed_render(torch, c2, stream, barrier, 7, 64, 1); c2.close()
                return {"value": st2["samples"] / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "spp": 64, "collisions_per_sample": st2["collisions"] / max(st2["samples"], 1),
                        "collisions_per_s": st2["collisions"] / (ms * 1e-3), "kernel": "wavefront volpathmis (lr_wavefront_mis.cuh: k_mis_track / shade_a / nee / shade_b over queues, TRACE stage; 128 B of p / f state per path)", "launches": st2.get("launches"),
                        "note": f"C3 frame under volpathmis, use_spectral_mis = {str(spectral).lower()} (volpathmis.cpp:79)"}
            extras["volpathmis_c3"] = mis(True)
            extras["volpathmis_c3_plain_mis"] = mis(False)
            extras["biovolpath_liver_single_mesh"] = bio(_sc.liver_single_mesh_bio, "Liver-SingleMesh as the fork ships it: biovolpath, `liver` medium, bumpmap over dielectric 1.38, "
                                                         "envmap, independent sampler (scenes/Liver-SingleMesh/mitsuba3/scene.xml)", 3.30e6)
            extras["biovolpath_liver_multi_mesh"] = bio(_sc.liver_multi_mesh_bio, "Liver-MultiMesh as its scene file lays it out: capsule shell with `glissonCapsule`, organ with "
                                                        "`parenchyma`, ldsampler (scenes/Liver-MultiMesh/mitsuba3/scene.xml)", 11.9e6)
    return dict(extras=extras, weak=weak, torch=torch, dist=dist, mi=mi, multi=multi, ctx=ctx, sc=sc, dev=dev, stream=stream, world=world, rank=rank, local=local,
                spp=spp, name=name, W=W, H=H, run=main_run, table=table, rooflines=rooflines, roofline=roofline, barrier=barrier, flush=flush, hbm_peak=hbm_peak,
                set_partition=set_partition, film=film)


def run_e2e(R, args):
    """Same render through the host API with HOST buffers: density grid pinned-host -> device (params.update():
    lr_grid_update recomputes majorant + super-grid), render of this rank's share, film all-reduce, developed image device ->
    pinned host on rank 0 (the other ranks read one pixel of their own copy)."""
    torch, ctx, sc, dev, world, rank = R["torch"], R["ctx"], R["sc"], R["dev"], R["world"], R["rank"]
    dist = R["dist"]; W, H, spp = R["W"], R["H"], R["spp"]
    has_grid = len(sc.grid_arrays) > 0                            # C1/C2 have no gridvolume: the step's only host traffic is the image
    gi = 0; n = ctx.grid_size(gi) if has_grid else 0
    grid_host = torch.from_numpy(np.ascontiguousarray(sc.grid_arrays[gi]).reshape(-1)).pin_memory() if has_grid else None
    out_host = torch.empty((H, W, sc.out_channels), dtype=torch.float32).pin_memory()
    steps = max(1, min(args.steps, 2))
    R["set_partition"](args.partition)

    def step(k):
        if has_grid:
            ctx.grid_update(gi, grid_host)                      # H2D of the step's input + majorant rebuild
        if world == 1:
            ctx.render(seed=1000 * k, spp=spp, out=out_host)     # render + develop + D2H
        else:
            ctx.film_clear(); ctx.render_lanes(1000 * k + (rank if args.partition == "seeds" else 0), spp, 0, 0)
            dist.all_reduce(R["film"]); ctx.film_develop(out_host)
        return float(out_host[H // 2, W // 2, 0])                # the result is read on the host

    step(-1)
    R["barrier"](); t0 = time.perf_counter()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True); e0.record(R["stream"])
    for k in range(steps):
        step(k)
    e1.record(R["stream"]); R["barrier"](); wall = time.perf_counter() - t0
    R["set_partition"]("seeds")
    t = torch.tensor([max(e0.elapsed_time(e1), 0.0)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()); frames = world if args.partition == "seeds" else 1
    return {"value": W * H * spp * steps * frames / (ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(n * 4),
            "d2h_bytes_per_step": int(out_host.numel() * 4), "steps": steps, "ms_per_step": ms / steps, "wall_ms_per_step": wall / steps * 1e3,
            "api": "params.update() [lr_grid_update, pinned host grid] + mi.render [lr_render -> pinned host image]"
                   + ("; N > 1: lr_set_lane_map share + in-place film all-reduce + lr_film_develop -> pinned host" if world > 1 else "")}


def run_prb(R, args):
    """C5: SphereLiverConstEnv inverse step = primal render + prbvolpath adjoint (spp_grad 64) + NCCL all-reduce of the
    128^3 density-grid gradient.  Partition "samples": ONE step, every rank takes spp / N samples of every pixel for the primal
    and for the adjoint (multi.render_distributed / backward_distributed), both reductions in place on the backend's buffers."""
    torch, mi, multi, dev, world, rank, dist = R["torch"], R["mi"], R["multi"], R["dev"], R["world"], R["rank"], R["dist"]
    from liverrenderer_b200 import scenes
    res, spp, gr = 512, 64, 128
    sc = mi.load_dict(scenes.sphere_liver_inverse(res=res, spp=spp, grid_res=gr, tracking=args.tracking))
    ctx = mi.Context(R["local"]).upload(sc); ctx.set_stream(R["stream"].cuda_stream)
    img = torch.empty((res, res, 3), dtype=torch.float32, device=dev)
    target = torch.full_like(img, 0.5)
    steps = max(1, min(args.steps, 3)); tot = {"samples": 0, "rays": 0, "collisions": 0, "ms_bwd": 0.0}
    strong = args.partition == "samples"; mode = "samples" if strong else "seeds"
    red = True if world > 1 else None

    def step(k):
        seed = k if strong else k + rank
        multi.render_distributed(ctx, seed, spp, rank, world, mode=mode, film_buf=red, out=img)
        g = (2.0 / img.numel()) * (img - target)                 # L2 loss adjoint (notebook cell 11); torch elementwise = plumbing
        if strong:
            grad = multi.backward_distributed(ctx, g, mi.sample_tea_32(seed, 1)[0], spp, rank, world, grad_buf=True, mode="samples")   # True: the backend's buffer, reduced in place
        else:
            ctx.grid_grad_zero(0); ctx.render_backward(g, seed=mi.sample_tea_32(seed, 1)[0], spp=spp)
            grad = multi.grid_grad_tensor(ctx, 0, dev)
            if world > 1:
                dist.all_reduce(grad)
        return ctx.stats(), grad

    step(-1); R["barrier"]()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True); e0.record(R["stream"])
    for k in range(steps):
        st, grad = step(k)
        tot["samples"] += st["samples"]; tot["rays"] += st["rays"]; tot["collisions"] += st["collisions"]; tot["ms_bwd"] += st["device_ms"]
    e1.record(R["stream"]); R["barrier"]()
    if not torch.is_tensor(grad):
        grad = torch.from_numpy(np.asarray(grad)).to(dev)
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    cnt = torch.tensor([tot["collisions"]], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX); dist.all_reduce(cnt)
    ms = float(t.item()); n = res * res * spp * steps * (1 if strong else world)
    ach = tot["collisions"] * float(PRB_BYTES_PER_COLLISION) / (tot["ms_bwd"] * 1e-3) / 1e9 if tot["ms_bwd"] > 0 else 0.0
    out = {"workload": f"C5 SphereLiverConstEnv prbvolpath {res}x{res}, spp {spp} primal + spp_grad {spp}, sigma_t grid {gr}^3",
           "grad_samples_per_s": n / (ms * 1e-3), "ms_per_step": ms / steps, "steps": steps, "adjoint_kernel_ms": tot["ms_bwd"] / steps, "scaling": "strong" if strong else "weak",
           "adjoint_collisions_per_s": tot["collisions"] / (tot["ms_bwd"] * 1e-3) if tot["ms_bwd"] > 0 else 0.0,
           "gradient_atomics_per_s": 8.0 * tot["collisions"] / (tot["ms_bwd"] * 1e-3) if tot["ms_bwd"] > 0 else 0.0,
           "roofline": {"bound": "hbm", "achieved": ach, "peak": R["hbm_peak"], "unit": "GB/s", "frac": ach / R["hbm_peak"], "traffic": None,
                        "kernel": "adjoint pass (k_wfp_* kernels, rank 0)", "algorithmic_bytes_per_unit": PRB_BYTES_PER_COLLISION,
                        "note": "64 B per adjoint collision (tap + 8 gradient atomics, pre-aggregation) over the adjoint's device time"},
           "allreduce_bytes": int(grad.numel() * 4) if world > 1 else 0, "grad_finite": bool(torch.isfinite(grad).all().item()),
           "grad_abs_max": float(grad.abs().max().item())}
    ctx.close()
    return out


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args); return
    R = run_b200(args)
    e2e = run_e2e(R, args)
    prb = None if args.no_prb else run_prb(R, args)
    cpu = None
    if R["rank"] == 0 and R["world"] == 1 and not args.no_cpu:
        _, cpu = cpu_sample(args, args.cpu_spp)
        if prb is not None:
            prb["cpu_baseline"] = cpu_sample_prb(args)
    if R["world"] > 1:
        R["dist"].barrier()
    if R["rank"] == 0:
        run = R["run"]; secs = run["ms"] * 1e-3; value = run["samples"] / secs; strong = args.partition == "samples"
        ex = R["extras"]; same_alg = None
        if cpu and ex.get("reference_order_tracking"):
            same_alg = {"value": ex["reference_order_tracking"]["value"] / cpu["value"],
                        "note": "same algorithm on both sides: GPU reference-order tracking (global majorant, the reference's draw order) / CPU oracle"}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": R["world"], "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": run["ms"] / args.steps, "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic",
                "config": {"workload": R["name"], "spp": R["spp"], "samples_per_step": run["samples"] / args.steps, "tracking": args.tracking,
                           "l2": "flushed between timed iterations (512 MiB write)",
                           "partition": ("one frame split by sample index: rank r renders spp/N samples of every pixel; in-place film all-reduce" if strong
                                         else "every rank renders the config with seed+rank; in-place film all-reduce"),
                           "same_config_as_reference_arm": False,
                           "reference_arm_note": "--impl reference renders the same scene at --cpu-spp samples per pixel with the reference's global-majorant tracking "
                                                 "(samples/s does not depend on spp); see same_algorithm_ratio for the like-for-like ratio"},
                "mrays_per_s": run["rays"] / secs / 1e6, "collisions_per_s": run["colls"] / secs,
                "collisions_per_sample": run["colls"] / max(run["samples"], 1), "rays_per_sample": run["rays"] / max(run["samples"], 1),
                "e2e": e2e, "gpu_launches": int(run["launches"]), "clocks": run["clk"], "roofline": R["roofline"], "rooflines": R["rooflines"],
                "stages": R["table"], "cpu_baseline": cpu, "same_algorithm_ratio": same_alg, "prb": prb, "weak": R["weak"]}
        line.update(ex)
        emit(line)
    if R["world"] > 1:
        R["dist"].destroy_process_group()


if __name__ == "__main__":
    main()
