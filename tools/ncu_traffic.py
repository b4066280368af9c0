"""DRAM traffic per launch of the tracking kernels from an ncu CSV
(ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --csv --log-file <csv> ...).
Writes the JSON bench.py reads for roofline.traffic.   usage: python tools/ncu_traffic.py <csv> <out.json> "<command>" """
import collections, csv, json, sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = None; per = collections.defaultdict(lambda: collections.defaultdict(float)); names = {}
for r in rows:
    if r[0] == "ID": hdr = r; continue
    if hdr is None: continue
    d = dict(zip(hdr, r))
    try: v = float(d["Metric Value"].replace(",", ""))
    except Exception: continue
    u = d["Metric Unit"]; scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0}.get(u, 1)
    per[d["ID"]][d["Metric Name"]] = v * scale; names[d["ID"]] = d["Kernel Name"]
sel = [i for i, n in names.items() if "k_wf_track" in n or "k_wf_nee" in n]
tot_b = sum(per[i]["dram__bytes_read.sum"] + per[i]["dram__bytes_write.sum"] for i in sel)
tot_ms = sum(per[i]["gpu__time_duration.sum"] for i in sel)
out = {"kernels": "k_wf_track_dda + k_wf_nee", "launches": len(sel), "traffic_bytes_per_launch": tot_b / max(len(sel), 1),
       "dram_read_bytes": sum(per[i]["dram__bytes_read.sum"] for i in sel), "dram_write_bytes": sum(per[i]["dram__bytes_write.sum"] for i in sel),
       "kernel_ms_under_ncu": tot_ms, "dram_GBps_under_ncu": tot_b / max(tot_ms, 1e-9) / 1e6, "command": sys.argv[3] if len(sys.argv) > 3 else ""}
json.dump(out, open(sys.argv[2], "w"), indent=1); print(json.dumps(out))
