# closing run of the round on the final build: C5 with both grid settings, whole GPU suite, smoke, bench line
set -x
mkdir -p gpurun_out
for g in 8 4; do LR_WF_GRID=$g timeout 300 python tools/prof_render.py c5 512 64 dda 3 2>&1 | tail -2 | cut -c1-100; done
for g in 8 4; do LR_WF_GRID=$g timeout 300 python tools/prof_render.py c2 1080 64 dda 2 2>&1 | tail -1 | cut -c1-100; done
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench rc=$?"
cut -c1-400 gpurun_out/bench_final.json
