#!/bin/bash
# time-sliced TRACE x occupancy: trace_min sweep on all four workloads
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
V=$GRAFT_REPO_ROOT/ray-tracing-practice_b200/build
O=gpurun_out/k_ab.txt; : > $O
for v in v_sl256x4 v_sl192x5 v_sl128x7 v_sl288x3; do
  export RTB_LIB=$V/$v.so
  for k in 6 8; do
    echo "== $v headline trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --reps 3 >> $O 2>&1
    echo "== $v shipped trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --workload shipped --reps 2 >> $O 2>&1
    echo "== $v 1m trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --workload spheres1m_4k_64spp_d8 --reps 2 >> $O 2>&1
    echo "== $v 10k trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --workload spheres10k_1080p_64spp_d16 --reps 3 >> $O 2>&1
  done
done
grep -E "^==|kernel_ms" $O | sed 's/"segments.*//' | sed 's/{"workload".*"kernel_ms"/ms/' | paste - - | cut -c1-120
