#!/bin/bash
# f3 timing on one GPU: drop-in executable (pipelined / serial) over the 100-frame shipped config, next to the reference program's own gpu_render loop
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 2400 python tools/time_frames.py --frames 100 --gpus 1 --ref-program 100 --ref > gpurun_out/g_frames_1gpu.jsonl 2> gpurun_out/g_frames_1gpu.err; echo "rc=$?"
cut -c1-400 gpurun_out/g_frames_1gpu.jsonl; tail -3 gpurun_out/g_frames_1gpu.err
