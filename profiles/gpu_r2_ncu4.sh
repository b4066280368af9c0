#!/bin/bash
# ncu evidence of the per-lane kernels behind SURVEY.md 8(f) ranks 2 and 4: k_render<biovolpath> on the fork's Liver-SingleMesh scene and
# k_render<volpathmis> on the C3 frame.  Launch list (time, DRAM / L2 bytes) + one --set full capture each.  Each command first runs plain.
set -x
O=gpurun_out; T=/tmp/ncu_r2; mkdir -p $O $T
for W in "single 1080 64" "mis 512 16"; do
  TAG=$(echo $W | cut -d' ' -f1); CMD="python tools/prof_bio.py $W"
  $CMD > $O/r2_plain_$TAG.log 2>&1 || exit 1
  ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none -c 200 --csv --log-file $O/r2_launches_$TAG.csv $CMD > $O/r2_ncu_l_$TAG.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:k_render -c 1 -f -o $T/$TAG $CMD > $O/r2_ncu_f_$TAG.log 2>&1
  ncu -i $T/$TAG.ncu-rep --page raw --csv > $O/r2_full_${TAG}_raw.csv
  ncu -i $T/$TAG.ncu-rep --page details --csv > $O/r2_full_${TAG}_details.csv
  SAMPLES=$(python -c "a='$W'.split(); print(int(a[1])**2*int(a[2]))")
  python tools/ncu_kernel_metrics.py $O/r2_launches_$TAG.csv $O/r2_full_${TAG}_raw.csv $SAMPLES $O/r2_kernel_metrics_$TAG.json "$CMD" > $O/r2_kernel_metrics_$TAG.txt 2>&1
  cat $O/r2_kernel_metrics_$TAG.txt
done
ls -la $O | tail -20
