#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/plane.jsonl
run() { echo -n "{\"variant\": \"$1\", \"workload\": \"$3\", \"r\": " >> gpurun_out/plane.jsonl; RTB_LIB=$2 timeout 300 python tools/prof_frame.py --reps 3 --workload $3 $4 | tr -d '\n' >> gpurun_out/plane.jsonl 2>>gpurun_out/plane.err; echo "}" >> gpurun_out/plane.jsonl; }
D=ray-tracing-practice_b200/librt_b200.so
N=ray-tracing-practice_b200/build/noplane.so
for i in 1 2; do
run fast $D shipped "--spp 625"
run off $N shipped "--spp 625"
run fast $D spheres100k_1080p_256spp_d32 ""
run off $N spheres100k_1080p_256spp_d32 ""
done
run fast $D spheres10k_1080p_64spp_d16 ""
run off $N spheres10k_1080p_64spp_d16 ""
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
