set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench1.json 2> gpurun_out/bench1.err; echo "bench rc=$?"
cat gpurun_out/bench1.json | cut -c1-600
timeout 300 python tools/prof_render.py c3 1080 64 dda 2 > gpurun_out/prof_plain.log 2>&1; cat gpurun_out/prof_plain.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_c3_dda.csv python tools/prof_render.py c3 1080 64 dda 1 > gpurun_out/ncu_l.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_ -s 60 -c 12 -o gpurun_out/c3_dda_full -f python tools/prof_render.py c3 1080 64 dda 1 > gpurun_out/ncu_f.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out
