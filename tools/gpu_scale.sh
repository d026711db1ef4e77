#!/bin/bash
# bench line at N GPUs of one box (N = first argument), headline workload; at N=8 also BASELINE configs[4]
N=$1
mkdir -p gpurun_out
run() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@"; }
run --steps 5 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "rc=$?" >> gpurun_out/bench_n$N.err
if [ "$N" = "8" ]; then
  run --steps 3 --warmup 3 --workload spheres100k_8k_1024spp_d32 > gpurun_out/bench_n8_8k.json 2> gpurun_out/bench_n8_8k.err; echo "rc=$?" >> gpurun_out/bench_n8_8k.err
fi
