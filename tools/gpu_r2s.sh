#!/bin/bash
# sphere grid: parity gates with the grid on (default), timing grid on / off on the four workloads
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out/s_ab.txt; : > $O
for g in 1 0; do
  echo "== grid=$g headline" >> $O; RTB_GRID=$g timeout 300 python tools/prof_frame.py --reps 3 --count >> $O 2>&1
  echo "== grid=$g 1m" >> $O; RTB_GRID=$g timeout 300 python tools/prof_frame.py --workload spheres1m_4k_64spp_d8 --reps 2 --count >> $O 2>&1
  echo "== grid=$g 10k" >> $O; RTB_GRID=$g timeout 300 python tools/prof_frame.py --workload spheres10k_1080p_64spp_d16 --reps 3 --count >> $O 2>&1
done
echo "== grid=1 shipped" >> $O; timeout 300 python tools/prof_frame.py --workload shipped --reps 2 >> $O 2>&1
timeout 1500 python -m pytest tests -m gpu -q -x -s > gpurun_out/s_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/s_pytest.log | cut -c1-300
grep -E "^==|kernel_ms" $O | sed 's/{"workload".*"kernel_ms"/ms/; s/"mean".*"per_ray"/per_ray/' | paste - - | cut -c1-330
