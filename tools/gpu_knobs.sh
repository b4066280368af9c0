# host-side knobs of the wavefront pipeline on C3 (no rebuild): CTAs per SM of the grid-stride stage kernels, batches in flight
mkdir -p gpurun_out
run() { echo "=== grid=$1 grid128=$2 runs=$3"; LR_WF_GRID=$1 LR_WF_GRID128=$2 LR_WF_RUNS=$3 timeout 300 python tools/prof_render.py c3 1080 256 dda 3 2>&1 | tail -2 | cut -c1-90; }
run 8 8 3
run 4 8 3
run 3 8 3
run 2 8 3
run 4 4 3
run 4 6 3
run 4 8 4
run 4 8 2
run 6 8 3
run 4 12 3
run 8 8 3
