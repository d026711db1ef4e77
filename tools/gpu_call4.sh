#!/bin/bash
# 2-GPU call: NCCL tile gather through the C-ABI, driver --split tiles, bench at N=2, texture pin dump
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/c4_gpus.txt
timeout 900 python -m pytest tests -m gpu -q -s -k "tiles or split or thread or parity or format" > gpurun_out/c4_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/c4_pytest.log
timeout 300 python tools/dump_texpin.py > gpurun_out/c4_texpin.txt 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/c4_bench_n2.json 2> gpurun_out/c4_bench_n2.err; echo "bench2 rc=$?"
timeout 600 python bench.py --steps 5 --warmup 3 --no-baselines > gpurun_out/c4_bench_n1.json 2> gpurun_out/c4_bench_n1.err; echo "bench1 rc=$?"
tail -5 gpurun_out/c4_pytest.log; cat gpurun_out/c4_texpin.txt | tail -3
tail -c 1200 gpurun_out/c4_bench_n2.json; tail -5 gpurun_out/c4_bench_n2.err
