#!/bin/bash
# A/B of kernel variants built with `make variant` (build/*.so), same frame, best of 3; then the GPU test suite on the default build
mkdir -p gpurun_out
: > gpurun_out/variants.jsonl
for v in base maxl1 sqrt alt base maxl1; do
  echo -n "{\"variant\": \"$v\", \"r\": " >> gpurun_out/variants.jsonl
  RTB_LIB=ray-tracing-practice_b200/build/$v.so timeout 300 python tools/prof_frame.py --reps 3 >> gpurun_out/variants.jsonl 2>gpurun_out/variant_$v.err
  echo "}" >> gpurun_out/variants.jsonl
done
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
