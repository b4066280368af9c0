#!/bin/bash
# C3 (1080^2, 64 spp, production tracking) after the TRACK fusion: launch list with time / DRAM / L2 bytes of every launch and two --set full
# windows (first iteration of the first batch; a later iteration), merged into r2_kernel_metrics.json.  The command first runs plain and must exit 0.
set -x
O=gpurun_out; T=/tmp/ncu_r2; mkdir -p $O $T
CMD="python tools/prof_render.py c3 1080 64 dda 1"
$CMD > $O/r2_plain_c3.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none -c 4000 --csv --log-file $O/r2_launches_c3_64spp.csv $CMD > $O/r2_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_wv_ -s 1 -c 10 -f -o $T/w1 $CMD > $O/r2_ncu2a.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_wv_ -s 32 -c 10 -f -o $T/w2 $CMD > $O/r2_ncu2b.log 2>&1
ncu -i $T/w1.ncu-rep --page raw --csv > $O/r2_full_c3_iter1_raw.csv
ncu -i $T/w2.ncu-rep --page raw --csv > $O/r2_full_c3_iter2_raw.csv
python tools/ncu_kernel_metrics.py $O/r2_launches_c3_64spp.csv $O/r2_full_c3_iter1_raw.csv,$O/r2_full_c3_iter2_raw.csv 74649600 $O/r2_kernel_metrics.json "$CMD" > $O/r2_kernel_metrics.txt 2>&1
cat $O/r2_kernel_metrics.txt; cat $O/r2_plain_c3.log
