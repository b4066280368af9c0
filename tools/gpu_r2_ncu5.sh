#!/bin/bash
# ncu evidence of the biovolpath WAVEFRONT on the fork's Liver-SingleMesh scene (1080^2, 64 spp): launch list with time / DRAM / L2 bytes of
# every launch, and a --set full window over the first two iterations.  The command first runs plain and must exit 0.
set -x
O=gpurun_out; T=/tmp/ncu_r2; mkdir -p $O $T
CMD="python tools/prof_bio.py single 1080 64"
$CMD > $O/r2_plain_biowf.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none -c 2000 --csv --log-file $O/r2_launches_biowf.csv $CMD > $O/r2_ncu_l_biowf.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_wv_|k_bio_" -s 1 -c 14 -f -o $T/biowf $CMD > $O/r2_ncu_f_biowf.log 2>&1
ncu -i $T/biowf.ncu-rep --page raw --csv > $O/r2_full_biowf_raw.csv
python tools/ncu_kernel_metrics.py $O/r2_launches_biowf.csv $O/r2_full_biowf_raw.csv 74649600 $O/r2_kernel_metrics_biowf.json "$CMD" > $O/r2_kernel_metrics_biowf.txt 2>&1
cat $O/r2_kernel_metrics_biowf.txt; cat $O/r2_plain_biowf.log
ls -la $O | tail -8
