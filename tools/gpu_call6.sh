#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/c6_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/c6_pytest.log
for q in 1 37 74 148 256; do
  echo "== headline queues=$q" >> gpurun_out/c6_ab.txt; RTB_QUEUES=$q timeout 300 python tools/prof_frame.py --reps 3 >> gpurun_out/c6_ab.txt 2>&1
done
for q in 1 148; do
  echo "== shipped queues=$q" >> gpurun_out/c6_ab.txt; RTB_QUEUES=$q timeout 300 python tools/prof_frame.py --workload shipped --reps 2 >> gpurun_out/c6_ab.txt 2>&1
  echo "== 1m queues=$q" >> gpurun_out/c6_ab.txt; RTB_QUEUES=$q timeout 300 python tools/prof_frame.py --workload spheres1m_4k_64spp_d8 --reps 2 >> gpurun_out/c6_ab.txt 2>&1
  echo "== 10k queues=$q" >> gpurun_out/c6_ab.txt; RTB_QUEUES=$q timeout 300 python tools/prof_frame.py --workload spheres10k_1080p_64spp_d16 --reps 3 >> gpurun_out/c6_ab.txt 2>&1
done
tail -3 gpurun_out/c6_pytest.log
grep -E "^==|kernel_ms" gpurun_out/c6_ab.txt | sed 's/"segments.*//' | cut -c1-170
