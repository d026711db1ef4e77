#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for cfg in "148 0" "444 1" "222 1" "888 1"; do set -- $cfg
  echo "== headline queues=$1 per_cta=$2" >> gpurun_out/c7_ab.txt; RTB_QUEUES=$1 RTB_QUEUE_PER_CTA=$2 timeout 300 python tools/prof_frame.py --reps 3 >> gpurun_out/c7_ab.txt 2>&1
  echo "== 1m queues=$1 per_cta=$2" >> gpurun_out/c7_ab.txt; RTB_QUEUES=$1 RTB_QUEUE_PER_CTA=$2 timeout 300 python tools/prof_frame.py --workload spheres1m_4k_64spp_d8 --reps 2 >> gpurun_out/c7_ab.txt 2>&1
  echo "== shipped queues=$1 per_cta=$2" >> gpurun_out/c7_ab.txt; RTB_QUEUES=$1 RTB_QUEUE_PER_CTA=$2 timeout 300 python tools/prof_frame.py --workload shipped --reps 2 >> gpurun_out/c7_ab.txt 2>&1
done
for c in 1 2 4; do
  echo "== headline queues=148 chunk=$c" >> gpurun_out/c7_ab.txt; RTB_SAMPLE_CHUNK=$c timeout 300 python tools/prof_frame.py --reps 3 >> gpurun_out/c7_ab.txt 2>&1
done
grep -E "^==|kernel_ms" gpurun_out/c7_ab.txt | sed 's/"segments.*//' | cut -c1-170
