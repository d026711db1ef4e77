#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -s -k "tiles or split or thread or parity or format or texture_unit or full_size" > gpurun_out/c5_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/c5_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/c5_bench_n2.json 2> gpurun_out/c5_bench_n2.err; echo "bench2 rc=$?"
timeout 1500 python tools/time_frames.py --frames 100 --ref > gpurun_out/c5_frames.jsonl 2> gpurun_out/c5_frames.err; echo "frames rc=$?"
tail -4 gpurun_out/c5_pytest.log; cut -c1-330 gpurun_out/c5_frames.jsonl
