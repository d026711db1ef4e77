#!/bin/bash
# every BASELINE config on one GPU with the current build; launch list of the device BVH build; bench line with the bvh_build leg
set -x
mkdir -p gpurun_out
timeout 900 python tools/run_configs.py > gpurun_out/all_configs.jsonl 2> gpurun_out/all_configs.err
timeout 300 python tools/prof_frame.py --builder gpu --spp 1 --reps 1 > gpurun_out/bvh_gpu_frame.json 2>&1 && \
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/bvh_gpu_launches.csv python tools/prof_frame.py --builder gpu --spp 1 --reps 1 > gpurun_out/bvh_gpu_ncu.log 2>&1
timeout 900 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?" >> gpurun_out/bench_n1.err
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
