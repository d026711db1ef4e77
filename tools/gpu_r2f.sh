#!/bin/bash
# per-ray work of primary rays alone; pooled kernel ncu with 48 parked hits, regroup off / on
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out/f_counts.txt; : > $O
for d in 1 2 3 32; do echo "== headline depth=$d" >> $O; timeout 300 python tools/prof_frame.py --depth $d --spp 64 --reps 2 --count >> $O 2>&1; done
for d in 1 2 50; do echo "== shipped depth=$d" >> $O; timeout 300 python tools/prof_frame.py --workload shipped --depth $d --spp 256 --reps 2 --count >> $O 2>&1; done
for r in 0 1; do
  CMD="python tools/prof_frame.py --kernel pool --spp 8 --reps 1"
  RTB_PARK_MIN=48 RTB_POOL_REGROUP=$r timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_pool -c 1 -f -o /tmp/pool_r$r $CMD > gpurun_out/f_ncu_pool_r$r.log 2>&1
  python tools/ncu_summary.py /tmp/pool_r$r.ncu-rep > gpurun_out/f_ncu_pool_park48_regroup${r}_summary.txt 2>&1
  python tools/ncu_by_line.py /tmp/pool_r$r.ncu-rep ray-tracing-practice_b200/librt_b200.so render_poolILi0ELb0ELb0E --top 400 > gpurun_out/f_ncu_pool_park48_regroup${r}_by_line.txt 2>&1
done
grep -E "^==|kernel_ms" $O | cut -c1-400
