# round-1 re-check after the bitmap/bumpmap change: new GPU tests first, then the whole GPU suite, smoke, the bench line
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "bump or library" > gpurun_out/pytest_bump.log 2>&1; echo "pytest bump rc=$?" >> gpurun_out/pytest_bump.log
tail -30 gpurun_out/pytest_bump.log
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log; tail -8 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench1.json 2> gpurun_out/bench1.err; echo "bench rc=$?"
cut -c1-1500 gpurun_out/bench1.json; tail -5 gpurun_out/bench1.err
