#!/bin/bash
# ncu evidence of the volpathmis WAVEFRONT on the C3 frame (512^2, 16 spp, spectral MIS, reference-order tracking): launch list with time / DRAM /
# L2 bytes of every launch, and a --set full window over the second and third iteration.  The command first runs plain and must exit 0.
set -x
O=gpurun_out; T=/tmp/ncu_r2; mkdir -p $O $T
CMD="python tools/prof_bio.py mis 512 16"
$CMD > $O/r2_plain_miswf.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none -c 3000 --csv --log-file $O/r2_launches_miswf.csv $CMD > $O/r2_ncu_l_miswf.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_wv_|k_mis_" -s 12 -c 24 -f -o $T/miswf $CMD > $O/r2_ncu_f_miswf.log 2>&1
ncu -i $T/miswf.ncu-rep --page raw --csv > $O/r2_full_miswf_raw.csv
python tools/ncu_kernel_metrics.py $O/r2_launches_miswf.csv $O/r2_full_miswf_raw.csv 4194304 $O/r2_kernel_metrics_miswf.json "$CMD" > $O/r2_kernel_metrics_miswf.txt 2>&1
cat $O/r2_kernel_metrics_miswf.txt; cat $O/r2_plain_miswf.log
