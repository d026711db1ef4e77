#!/bin/bash
# hoisted floor quad A/B (runtime switch RTB_HOIST) on the headline frame and the shipped scene, then the GPU suite
mkdir -p gpurun_out
: > gpurun_out/hoist.jsonl
for h in 1 0 1 0; do
  echo -n "{\"hoist\": $h, \"r\": " >> gpurun_out/hoist.jsonl
  RTB_HOIST=$h timeout 300 python tools/prof_frame.py --reps 3 --count | tr -d '\n' >> gpurun_out/hoist.jsonl 2>gpurun_out/hoist_$h.err
  echo "}" >> gpurun_out/hoist.jsonl
done
RTB_HOIST=1 timeout 600 python tools/run_configs.py shipped spheres10k_1080p_64spp_d16 spheres1m_4k_64spp_d8 > gpurun_out/hoist1_configs.jsonl 2>&1
RTB_HOIST=0 timeout 600 python tools/run_configs.py shipped spheres10k_1080p_64spp_d16 spheres1m_4k_64spp_d8 > gpurun_out/hoist0_configs.jsonl 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
