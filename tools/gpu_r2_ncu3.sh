#!/bin/bash
# ncu evidence of the prbvolpath wavefront (C5 backward = primal pass + adjoint replay): launch list with DRAM / L2 bytes of every launch and a
# --set full window over the first two iterations of the ADJOINT pass.  The command first runs plain and must exit 0.
set -x
O=gpurun_out; T=/tmp/ncu_r2; mkdir -p $O $T
CMD5="python tools/prof_render.py c5 512 64 dda 1"
$CMD5 > $O/r2_plain_c5.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none -c 4000 --csv --log-file $O/r2_launches_c5.csv $CMD5 > $O/r2_ncu5.log 2>&1
python - <<'PY'
import csv, re
rows=[r for r in csv.reader(open("gpurun_out/r2_launches_c5.csv")) if len(r)>5]
hdr=None; order=[]; seen=set()
for r in rows:
    if r[0]=="ID": hdr=r; continue
    if hdr is None: continue
    d=dict(zip(hdr,r))
    if d["ID"] not in seen and "k_w" in d["Kernel Name"]: seen.add(d["ID"]); order.append(d["Kernel Name"])
# index (among k_w* launches) of the second k_wvp_gen = start of the adjoint pass of the backward call (the first render is the primal image)
gens=[i for i,n in enumerate(order) if "k_wvp_gen" in n]
print("k_w launches", len(order), "gen at", gens)
open("/tmp/ncu_r2/adj_start","w").write(str(gens[-1]))
PY
S=$(cat $T/adj_start)
ncu --set full --clock-control none --import-source on -k regex:k_w -s $S -c 26 -f -o $T/adj $CMD5 > $O/r2_ncu6.log 2>&1
ncu -i $T/adj.ncu-rep --page raw --csv > $O/r2_full_c5_adjoint_raw.csv
python tools/ncu_kernel_metrics.py $O/r2_launches_c5.csv $O/r2_full_c5_adjoint_raw.csv 16777216 $O/r2_kernel_metrics_c5.json "$CMD5" > $O/r2_kernel_metrics_c5.txt 2>&1
cat $O/r2_kernel_metrics_c5.txt
