#!/bin/bash
# code-size variants: out-of-line plane fallback
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
V=$GRAFT_REPO_ROOT/ray-tracing-practice_b200/build
O=gpurun_out/w_ab.txt; : > $O
for v in default $VARIANTS; do
  if [ $v = default ]; then unset RTB_LIB; else export RTB_LIB=$V/$v.so; fi
  echo "== $v headline" >> $O; timeout 300 python tools/prof_frame.py --reps 3 >> $O 2>&1
  echo "== $v shipped" >> $O; timeout 300 python tools/prof_frame.py --workload shipped --reps 2 >> $O 2>&1
  echo "== $v 1m" >> $O; timeout 300 python tools/prof_frame.py --workload spheres1m_4k_64spp_d8 --reps 2 >> $O 2>&1
  echo "== $v 10k" >> $O; timeout 300 python tools/prof_frame.py --workload spheres10k_1080p_64spp_d16 --reps 3 >> $O 2>&1
done
grep -E "^==|kernel_ms" $O | sed 's/"segments.*//' | sed 's/{"workload".*"kernel_ms"/ms/' | paste - - | cut -c1-100
