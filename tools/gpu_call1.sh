#!/bin/bash
# first GPU call of round 2: parity gates, bench, per-config baselines
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/c1_gpus.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/c1_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/c1_pytest.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/c1_bench.json 2> gpurun_out/c1_bench.err; echo "bench rc=$?"
timeout 900 python tools/run_configs.py --ref > gpurun_out/c1_all_configs_ref.jsonl 2> gpurun_out/c1_all_configs.err; echo "configs rc=$?"
tail -c 1500 gpurun_out/c1_pytest.log
tail -c 600 gpurun_out/c1_bench.json
