#!/bin/bash
# sign-of-reciprocal octant indices (v_sign) and the single-pop-site loop (v_pop1) against the committed default
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
V=$GRAFT_REPO_ROOT/ray-tracing-practice_b200/build
O=gpurun_out/o_ab.txt; : > $O
for v in default v_un2 v_pf v_un2pf; do
  if [ $v = default ]; then unset RTB_LIB; else export RTB_LIB=$V/$v.so; fi
  for k in 6; do
    echo "== $v headline trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --reps 3 >> $O 2>&1
    echo "== $v shipped trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --workload shipped --reps 2 >> $O 2>&1
    echo "== $v 1m trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --workload spheres1m_4k_64spp_d8 --reps 2 >> $O 2>&1
    echo "== $v 10k trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --workload spheres10k_1080p_64spp_d16 --reps 3 >> $O 2>&1
  done
done
export RTB_LIB=$V/v_un2.so
for k in 4 8; do
  echo "== v_un2 headline trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --reps 3 >> $O 2>&1
  echo "== v_un2 shipped trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --workload shipped --reps 2 >> $O 2>&1
done
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/o_pytest_un2pf.log 2>&1; echo "pytest pop1 rc=$?"
grep -E "^==|kernel_ms" $O | sed 's/"segments.*//' | sed 's/{"workload".*"kernel_ms"/ms/' | paste - - | cut -c1-120
