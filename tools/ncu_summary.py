"""Summarise ncu outputs for profiles/: (1) a launch list CSV (--metrics gpu__time_duration.sum) -> per-kernel totals and shares;
(2) a --set full report (.ncu-rep, read with `ncu -i ... --page raw --csv`) -> the handful of metrics the roofline needs.
usage: python tools/ncu_summary.py launches <csv>   |   python tools/ncu_summary.py full <ncu-rep>"""
import collections, csv, subprocess, sys, io

def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = None; agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows:
        if r[0] == "ID": hdr = r; continue
        if hdr is None: continue
        d = dict(zip(hdr, r))
        try: v = float(d["Metric Value"].replace(",", ""))
        except Exception: continue
        u = d["Metric Unit"]; v = v / 1e3 if u in ("ns", "nsecond") else (v * 1e3 if u in ("ms", "msecond") else v)
        name = d["Kernel Name"].split("(")[0].replace("void ", "")
        agg[name][0] += 1; agg[name][1] += v
    tot = sum(v[1] for v in agg.values())
    print(f"{'kernel':58s} {'launches':>8s} {'total ms':>10s} {'share':>7s} {'avg us':>9s}")
    for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"{k[:58]:58s} {v[0]:8d} {v[1] / 1e3:10.3f} {v[1] / tot * 100:6.1f}% {v[1] / v[0]:9.1f}")
    print(f"{'TOTAL':58s} {sum(v[0] for v in agg.values()):8d} {tot / 1e3:10.3f}")

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum", "l1tex__t_bytes.sum",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio", "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio"]

def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out))); hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    for vals in rows[2:]:
        print("==", vals[ki][:90])
        for i, h in enumerate(hdr):
            if h in WANT: print(f"   {h:86s} {vals[i]:>16s} {units[i]}")

if __name__ == "__main__":
    (launches if sys.argv[1] == "launches" else full)(sys.argv[2])
