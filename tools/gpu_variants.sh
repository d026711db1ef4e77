#!/bin/bash
# A/B of kernel variants built with `make variant` (build/*.so) against the default library, same frame, best of 3
mkdir -p gpurun_out
: > gpurun_out/variants.jsonl
run() { # name, lib, extra args
  echo -n "{\"variant\": \"$1\", \"r\": " >> gpurun_out/variants.jsonl
  RTB_LIB=$2 timeout 300 python tools/prof_frame.py --reps 3 $3 | tr -d '\n' >> gpurun_out/variants.jsonl 2>gpurun_out/variant_$1.err
  echo "}" >> gpurun_out/variants.jsonl
}
D=ray-tracing-practice_b200/librt_b200.so
run base $D ""
run preload ray-tracing-practice_b200/build/preload.so ""
run chunk1 $D "--chunk 1"
run chunk4 $D "--chunk 4"
run base2 $D ""
run preload2 ray-tracing-practice_b200/build/preload.so ""
