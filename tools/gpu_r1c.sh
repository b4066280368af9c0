# round-1 closing evidence for the final build: plain run, ncu launch list (C3 and C3 with the bump-mapped boundary), one full capture of the top kernels
set -x
mkdir -p gpurun_out
timeout 300 python tools/prof_render.py c3 1080 64 dda 2 > gpurun_out/prof_plain.log 2>&1; cat gpurun_out/prof_plain.log
LR_SURFACE=bumpmap timeout 300 python tools/prof_render.py c3 1080 64 dda 2 > gpurun_out/prof_plain_bump.log 2>&1; cat gpurun_out/prof_plain_bump.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_c3_final.csv python tools/prof_render.py c3 1080 64 dda 1 > gpurun_out/ncu_l.log 2>&1; echo "ncu launches rc=$?"
LR_SURFACE=bumpmap timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_c3_bump.csv python tools/prof_render.py c3 1080 64 dda 1 > gpurun_out/ncu_lb.log 2>&1; echo "ncu launches bump rc=$?"
LR_SURFACE=bumpmap timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_shade -s 40 -c 8 -o gpurun_out/c3_bump_full -f python tools/prof_render.py c3 1080 64 dda 1 > gpurun_out/ncu_f.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out
