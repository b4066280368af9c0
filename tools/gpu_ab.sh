# A/B of backend builds: usage: bash tools/gpu_ab.sh <variant.so> [<variant.so> ...]   (paths relative to the repo root)
mkdir -p gpurun_out
for v in "$@"; do
  echo "=== $v"
  LR_B200_LIB=$PWD/$v timeout 300 python tools/prof_render.py c3 1080 64 dda 3 2>&1 | tail -2
  LR_B200_LIB=$PWD/$v timeout 300 python tools/prof_render.py c5 512 64 dda 2 2>&1 | tail -1
done
