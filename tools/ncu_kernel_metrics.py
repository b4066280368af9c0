"""profiles/r2_kernel_metrics.json -- what bench.py's `rooflines` quote from ncu, built from two captures of ONE command:
  (1) the launch list with DRAM / L2 bytes of EVERY launch of the render
      ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none --csv --log-file <csv> <command>
  (2) a `--set full` capture of a window of launches of the same command (issue-slot utilisation, threads per instruction)
      ncu --set full --clock-control none --import-source on -k regex:k_wv_ -s <skip> -c <count> -o <rep> <command>
usage: python tools/ncu_kernel_metrics.py <launches.csv> <full.ncu-rep or raw.csv>[,<more>...] <samples of the render> <out.json> "<command>" """
import collections, csv, io, json, re, subprocess, sys

csv_path, rep, samples, out_path, command = sys.argv[1], sys.argv[2], float(sys.argv[3]), sys.argv[4], sys.argv[5]


def kname(n):
    m = re.search(r"(k_w[vf]p?_\w+|k_\w+)(<[^>]*>)?", n)
    if not m:
        return n
    base, targs = m.group(1), m.group(2) or ""
    if base == "k_wv_trace":
        a = re.search(r"(\d+)", targs); return f"k_wv_trace<{a.group(1) if a else '?'}>"
    return base


SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0, "second": 1e3, "s": 1e3}
# ---- (1) launch list ----
rows = [r for r in csv.reader(open(csv_path)) if len(r) > 5]
hdr = None; per = collections.defaultdict(dict); names = {}
for r in rows:
    if r[0] == "ID":
        hdr = r; continue
    if hdr is None:
        continue
    d = dict(zip(hdr, r))
    try:
        v = float(d["Metric Value"].replace(",", ""))
    except Exception:
        continue
    per[d["ID"]][d["Metric Name"]] = v * SCALE.get(d["Metric Unit"], 1); names[d["ID"]] = kname(d["Kernel Name"])
K = collections.defaultdict(lambda: {"launches": 0, "ms_under_ncu": 0.0, "dram_bytes": 0.0, "l2_bytes": 0.0})
for i, n in names.items():
    k = K[n]; m = per[i]; k["launches"] += 1; k["ms_under_ncu"] += m.get("gpu__time_duration.sum", 0.0)
    k["dram_bytes"] += m.get("dram__bytes_read.sum", 0.0) + m.get("dram__bytes_write.sum", 0.0); k["l2_bytes"] += m.get("lts__t_bytes.sum", 0.0)
tot_ms = sum(k["ms_under_ncu"] for k in K.values()); tot_dram = sum(k["dram_bytes"] for k in K.values())
kernels = {}
for n, k in sorted(K.items(), key=lambda kv: -kv[1]["ms_under_ncu"]):
    kernels[n] = {"launches": k["launches"], "ms_under_ncu": round(k["ms_under_ncu"], 3), "share_of_gpu_time": round(k["ms_under_ncu"] / tot_ms, 4),
                  "dram_bytes_per_launch": k["dram_bytes"] / k["launches"], "dram_GBps": k["dram_bytes"] / max(k["ms_under_ncu"], 1e-9) / 1e6,
                  "l2_GBps": k["l2_bytes"] / max(k["ms_under_ncu"], 1e-9) / 1e6, "dram_bytes_per_sample": k["dram_bytes"] / samples}
# ---- (2) full-set window: duration-weighted means per kernel ----
acc = collections.defaultdict(lambda: collections.defaultdict(float)); wsum = collections.defaultdict(float); cnt = collections.Counter()
for one in rep.split(","):                       # each: an .ncu-rep, or its `--page raw --csv` export
    raw = open(one).read() if one.endswith(".csv") else subprocess.run(["ncu", "-i", one, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rr = list(csv.reader(io.StringIO(raw)))
    if len(rr) <= 2:
        continue
    h = rr[0]; ki = h.index("Kernel Name")
    want = {"issue_active_pct": "smsp__issue_active.avg.pct_of_peak_sustained_active", "threads_per_inst": "smsp__thread_inst_executed_per_inst_executed.ratio",
            "warps_active_pct": "sm__warps_active.avg.pct_of_peak_sustained_active", "registers": "launch__registers_per_thread",
            "l1_lsu_wavefronts_pct": "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
            "long_scoreboard_per_issue": "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "l2_hit_pct": "lts__t_sector_hit_rate.pct"}
    def num(x):
        try:
            return float(x.replace(",", ""))
        except ValueError:
            return None
    for v in rr[2:]:
        n = kname(v[ki]); w = num(v[h.index("gpu__time_duration.sum")]) or 0.0
        if w <= 0:
            continue
        wsum[n] += w; cnt[n] += 1
        for key, col in want.items():
            if col in h and num(v[h.index(col)]) is not None:
                acc[n][key] += w * num(v[h.index(col)])
for n in wsum:
    e = kernels.setdefault(n, {})
    for key in acc[n]:
        e[key] = round(acc[n][key] / wsum[n], 3)
    e["full_set_launches"] = cnt[n]
    if "issue_active_pct" in e and "threads_per_inst" in e:
        e["lane_issue_utilisation"] = round(e["issue_active_pct"] / 100.0 * e["threads_per_inst"] / 32.0, 4)
out = {"source": f"profiles/{out_path.split('/')[-1]} <- {csv_path.split('/')[-1]} + {rep.split('/')[-1]}", "command": command, "samples": samples,
       "gpu_ms_under_ncu": round(tot_ms, 2), "dram_bytes_per_sample": tot_dram / samples, "dram_bytes_total": tot_dram, "kernels": kernels,
       "note": "per-launch times under ncu are serialised and cold-cache: shares, bytes and ratios are the evidence, not the absolute times"}
json.dump(out, open(out_path, "w"), indent=1)
print(f"dram bytes / sample {out['dram_bytes_per_sample']:.1f}; gpu ms under ncu {tot_ms:.1f}")
for n, e in kernels.items():
    print(n.ljust(26), " ".join(f"{k}={e[k]}" for k in ("launches", "share_of_gpu_time", "issue_active_pct", "threads_per_inst", "lane_issue_utilisation", "l1_lsu_wavefronts_pct", "registers") if k in e),
          f"dram/launch {e.get('dram_bytes_per_launch', 0) / 1e6:.1f} MB  dram {e.get('dram_GBps', 0):.0f} GB/s  l2 {e.get('l2_GBps', 0):.0f} GB/s")
