#!/bin/bash
# every BASELINE config on one GPU with the reference CUDA kernel beside it (kept evidence: profiles/r02_all_configs_ref.jsonl)
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python tools/run_configs.py --ref > gpurun_out/c_all_configs_ref.jsonl 2> gpurun_out/c_all_configs_ref.err; echo "rc=$?"
timeout 600 python tools/prof_frame.py --workload shipped --reps 2 --count > gpurun_out/c_shipped_count.json 2>&1
cut -c1-420 gpurun_out/c_all_configs_ref.jsonl; tail -3 gpurun_out/c_all_configs_ref.err
