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
for v in $VARIANTS; do
  run base $D "$ARGS"
  run $v ray-tracing-practice_b200/build/$v.so "$ARGS"
done
