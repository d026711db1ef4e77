#!/bin/bash
# 8 GPUs of one box: headline bench line and BASELINE configs[4] (8K, 1024 spp) tile-partitioned with the NCCL gather
set -x
mkdir -p gpurun_out
run() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 "$@"; }
run --steps 5 --warmup 3 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; echo "rc=$?" >> gpurun_out/bench_n8.err
run --steps 3 --warmup 3 --workload spheres100k_8k_1024spp_d32 > gpurun_out/bench_n8_8k.json 2> gpurun_out/bench_n8_8k.err; echo "rc=$?" >> gpurun_out/bench_n8_8k.err
