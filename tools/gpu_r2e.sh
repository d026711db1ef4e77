#!/bin/bash
# smem-counter time slicing (shape 5) sweeps; tile rotation A/B by per-rank times; ncu of the pooled kernel with / without regroup
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
V=$GRAFT_REPO_ROOT/ray-tracing-practice_b200/build
O=gpurun_out/e_ab.txt; : > $O
export RTB_LIB=$V/v_slice5.so
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/e_pytest_slice5.log 2>&1; echo "pytest slice5 rc=$?"
for k in 1 4 8 12 16 20; do
  echo "== slice5 headline trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --reps 3 >> $O 2>&1
done
for k in 4 8 12 16; do
  echo "== slice5 shipped trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --workload shipped --reps 2 >> $O 2>&1
done
for k in 8; do
  echo "== slice5 1m trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --workload spheres1m_4k_64spp_d8 --reps 2 >> $O 2>&1
  echo "== slice5 10k trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --workload spheres10k_1080p_64spp_d16 --reps 3 >> $O 2>&1
done
export RTB_LIB=$V/v_slice.so
for k in 4 8; do
  echo "== slice4 shipped trace_min=$k" >> $O; RTB_TRACE_MIN=$k timeout 300 python tools/prof_frame.py --workload shipped --reps 2 >> $O 2>&1
done
unset RTB_LIB
echo "== default shipped" >> $O; timeout 300 python tools/prof_frame.py --workload shipped --reps 2 >> $O 2>&1
echo "---- kronecker rotation (default build)" > gpurun_out/e_rank_times.txt
for n in 8 4 2; do echo "-- $n ranks" >> gpurun_out/e_rank_times.txt; timeout 600 python tools/rank_times.py $n 256 >> gpurun_out/e_rank_times.txt 2>&1; done
echo "---- hashed rotation (round-2 start)" >> gpurun_out/e_rank_times.txt
for n in 8 2; do echo "-- $n ranks" >> gpurun_out/e_rank_times.txt; RTB_LIB=$V/v_rot0.so timeout 600 python tools/rank_times.py $n 256 >> gpurun_out/e_rank_times.txt 2>&1; done
# ncu: pooled kernel, regroup off / on (lanes per instruction in the shading code)
for r in 0 1; do
  CMD="python tools/prof_frame.py --kernel pool --spp 8 --reps 1"
  RTB_POOL_REGROUP=$r timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_pool -c 1 -f -o /tmp/pool_r$r $CMD > gpurun_out/e_ncu_pool_r$r.log 2>&1
  python tools/ncu_summary.py /tmp/pool_r$r.ncu-rep > gpurun_out/e_ncu_pool_regroup${r}_summary.txt 2>&1
  python tools/ncu_by_line.py /tmp/pool_r$r.ncu-rep ray-tracing-practice_b200/librt_b200.so render_poolILi0ELb0ELb0E --top 400 > gpurun_out/e_ncu_pool_regroup${r}_by_line.txt 2>&1
done
tail -3 gpurun_out/e_pytest_slice5.log
grep -E "^==|kernel_ms" $O | sed 's/"segments.*//' | sed 's/{"workload".*"kernel_ms"/ms/' | paste - - | cut -c1-120
cat gpurun_out/e_rank_times.txt | cut -c1-60
head -12 gpurun_out/e_ncu_pool_regroup1_by_line.txt
