"""One line per profiled launch of an .ncu-rep with the metrics that matter for these kernels.  usage: python tools/ncu_table.py <rep>"""
import csv, subprocess, sys, io
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out))); hdr = rows[0]
cols = [("gpu__time_duration.sum", "ms"), ("smsp__inst_executed.sum", "warp-inst"), ("smsp__thread_inst_executed_per_inst_executed.ratio", "thr/inst"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"), ("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "L1wave%"),
        ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "longsb"), ("dram__bytes_read.sum", "dramR MB"), ("dram__bytes_write.sum", "dramW MB"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"), ("lts__t_sector_hit_rate.pct", "L2hit%"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("launch__registers_per_thread", "regs")]
ki = hdr.index("Kernel Name")
print("kernel".ljust(30), " ".join(n.rjust(10) for _, n in cols))
for v in rows[2:]:
    vals = []
    for c, n in cols:
        x = v[hdr.index(c)] if c in hdr else ""
        try: x = f"{float(x.replace(',', '')):.4g}"
        except ValueError: pass
        vals.append(x[:10].rjust(10))
    print(v[ki].replace("void ", "").split("(")[0][:30].ljust(30), " ".join(vals))
