#!/bin/bash
# time-sliced TRACE variant: parity + timing sweep over trace_min; pooled megakernel with / without regroup by material; per-rank times of an 8-way split
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
V=$GRAFT_REPO_ROOT/ray-tracing-practice_b200/build
O=gpurun_out/d_ab.txt; : > $O
echo "== default headline" >> $O; timeout 300 python tools/prof_frame.py --reps 3 >> $O 2>&1
export RTB_LIB=$V/v_slice.so
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_bvh_build.py -m gpu -q -x > gpurun_out/d_pytest_slice.log 2>&1; echo "pytest slice rc=$?"
for k in 1 8 12 16 20 24 28; do
  echo "== slice headline trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --reps 3 >> $O 2>&1
done
for k in 12 20; do
  echo "== slice shipped trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --workload shipped --reps 2 >> $O 2>&1
  echo "== slice 1m trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --workload spheres1m_4k_64spp_d8 --reps 2 >> $O 2>&1
  echo "== slice 10k trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --workload spheres10k_1080p_64spp_d16 --reps 3 >> $O 2>&1
done
unset RTB_LIB
for cfg in "0 32" "1 32" "1 40" "1 48" "0 48"; do set -- $cfg
  echo "== pool headline64 regroup=$1 park_min=$2" >> $O; RTB_POOL_REGROUP=$1 RTB_PARK_MIN=$2 timeout 300 python tools/prof_frame.py --kernel pool --spp 64 --reps 3 >> $O 2>&1
  echo "== pool shipped256 regroup=$1 park_min=$2" >> $O; RTB_POOL_REGROUP=$1 RTB_PARK_MIN=$2 timeout 300 python tools/prof_frame.py --kernel pool --workload shipped --spp 256 --reps 3 >> $O 2>&1
done
echo "== mega headline64" >> $O; timeout 300 python tools/prof_frame.py --spp 64 --reps 3 >> $O 2>&1
echo "== mega shipped256" >> $O; timeout 300 python tools/prof_frame.py --workload shipped --spp 256 --reps 3 >> $O 2>&1
timeout 600 python tools/rank_times.py 8 256 > gpurun_out/d_rank_times.txt 2>&1
tail -3 gpurun_out/d_pytest_slice.log
grep -E "^==|kernel_ms" $O | sed 's/"segments.*//' | cut -c1-170
cat gpurun_out/d_rank_times.txt
