#!/bin/bash
# occupancy variants, second batch
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
V=$GRAFT_REPO_ROOT/ray-tracing-practice_b200/build
O=gpurun_out/i_ab.txt; : > $O
for v in v_setup_t224 v_t224 v_s192x5 v_s288x3; do
  export RTB_LIB=$V/$v.so
  echo "== $v headline" >> $O; timeout 300 python tools/prof_frame.py --reps 3 >> $O 2>&1
  echo "== $v shipped" >> $O; timeout 300 python tools/prof_frame.py --workload shipped --reps 2 >> $O 2>&1
  echo "== $v 1m" >> $O; timeout 300 python tools/prof_frame.py --workload spheres1m_4k_64spp_d8 --reps 2 >> $O 2>&1
  echo "== $v 10k" >> $O; timeout 300 python tools/prof_frame.py --workload spheres10k_1080p_64spp_d16 --reps 3 >> $O 2>&1
done
export RTB_LIB=$V/v_s224_slice.so
for k in 1 6 8 10; do
  echo "== v_s224_slice headline trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --reps 3 >> $O 2>&1
  echo "== v_s224_slice shipped trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --workload shipped --reps 2 >> $O 2>&1
done
grep -E "^==|kernel_ms" $O | sed 's/"segments.*//' | sed 's/{"workload".*"kernel_ms"/ms/' | paste - - | cut -c1-120
