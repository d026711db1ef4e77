#!/bin/bash
# final kernel: trace_min and sample-chunk sweeps
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out/v_ab.txt; : > $O
for k in 4 5 6 7 8 10; do
  echo "== headline trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --reps 3 >> $O 2>&1
  echo "== shipped trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --workload shipped --reps 2 >> $O 2>&1
  echo "== 1m trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --workload spheres1m_4k_64spp_d8 --reps 2 >> $O 2>&1
  echo "== 10k trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --workload spheres10k_1080p_64spp_d16 --reps 3 >> $O 2>&1
done
for c in 1 2 4 8; do
  echo "== headline chunk=$c" >> $O; RTB_SAMPLE_CHUNK=$c timeout 300 python tools/prof_frame.py --reps 3 >> $O 2>&1
  echo "== shipped chunk=$c" >> $O; RTB_SAMPLE_CHUNK=$c timeout 300 python tools/prof_frame.py --workload shipped --reps 2 >> $O 2>&1
done
grep -E "^==|kernel_ms" $O | sed 's/"segments.*//' | sed 's/{"workload".*"kernel_ms"/ms/' | paste - - | cut -c1-100
