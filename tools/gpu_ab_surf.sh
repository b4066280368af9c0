# A/B: register budgets of the stage kernels (LR_WF_OCC resident 256-thread blocks, LR_WF_OCC_SURF resident 128-thread blocks per SM)
mkdir -p gpurun_out
run() { echo "=== $1 grid=$2 grid128=$3"; LR_B200_LIB=$1 LR_WF_GRID=$2 LR_WF_GRID128=$3 timeout 300 python tools/prof_render.py c3 1080 128 dda 3 2>&1 | tail -2; }
B=$PWD/liverrenderer_b200/csrc
run $B/liblr_b200.so 8 8
run $B/ab/liblr_occ5.so 10 8
run $B/ab/liblr_occ5.so 5 8
run $B/ab/liblr_occ3.so 6 8
run $B/ab/liblr_occ3.so 9 8
run $B/ab/liblr_surf10.so 8 10
run $B/ab/liblr_surf10.so 8 20
run $B/liblr_b200.so 4 8
run $B/liblr_b200.so 12 8
