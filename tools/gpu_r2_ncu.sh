#!/bin/bash
# ncu evidence of round 2 (run on the GPU box through gpurun): the launch list + DRAM bytes of one C3 render, a --set full window of
# its stage kernels, the adjoint's kernels, and the launch list of the bench command.  Every command first runs plain and must exit 0.
# The .ncu-rep files stay on the box (too large to bring back): their raw / source pages are exported as CSV here.
set -x
O=gpurun_out; T=/tmp/ncu_r2; mkdir -p $O $T
CMD="python tools/prof_render.py c3 1080 64 dda 1"
$CMD > $O/r2_plain_c3.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none -c 4000 --csv --log-file $O/r2_launches_c3_64spp.csv $CMD > $O/r2_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_wv_ -s 13 -c 26 -f -o $T/r2_full_c3_64spp $CMD > $O/r2_ncu2.log 2>&1
ncu -i $T/r2_full_c3_64spp.ncu-rep --page raw --csv > $O/r2_full_c3_64spp_raw.csv
python tools/ncu_kernel_metrics.py $O/r2_launches_c3_64spp.csv $T/r2_full_c3_64spp.ncu-rep 74649600 $O/r2_kernel_metrics.json "$CMD" > $O/r2_kernel_metrics.txt 2>&1
for k in k_wv_track_dda k_wv_shade_a_surface k_wv_shade_b_surface k_wv_shade_b_medium k_wv_gen; do
  ncu -i $T/r2_full_c3_64spp.ncu-rep --page source --csv --print-source sass,cuda --kernel-name regex:$k --launch-skip 1 --launch-count 1 2>/dev/null | gzip > $O/r2_src_$k.csv.gz
done
CMD5="python tools/prof_render.py c5 512 64 dda 1"
$CMD5 > $O/r2_plain_c5.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:k_wfp_ -s 8 -c 24 -f -o $T/r2_full_c5 $CMD5 > $O/r2_ncu3.log 2>&1
ncu -i $T/r2_full_c5.ncu-rep --page raw --csv > $O/r2_full_c5_raw.csv
for k in k_wfp_track k_wfp_nee k_wfp_shade_surface; do
  ncu -i $T/r2_full_c5.ncu-rep --page source --csv --print-source sass,cuda --kernel-name regex:$k --launch-skip 1 --launch-count 1 2>/dev/null | gzip > $O/r2_src_$k.csv.gz
done
BCMD="python bench.py --steps 1 --warmup 3 --no-extras --no-prb --no-cpu"
$BCMD > $O/r2_bench_plain.json 2> $O/r2_bench_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 2400 --csv --log-file $O/r2_bench_launches.csv $BCMD > $O/r2_ncu4.log 2>&1
du -sh $O; ls -la $O
