# multi-GPU bench lines (run under gpurun --gpus 8): N = 8, 4, 2 through torchrun exactly as the driver launches them
mkdir -p gpurun_out
for n in 8 4 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n --steps 3 --warmup 3 > gpurun_out/bench_${n}gpu.json 2> gpurun_out/bench_${n}gpu.err
  echo "N=$n rc=$?"; tail -c 1500 gpurun_out/bench_${n}gpu.json | cut -c1-400; tail -3 gpurun_out/bench_${n}gpu.err
done
