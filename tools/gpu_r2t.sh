#!/bin/bash
# grid kernel: timing + ncu per-line profile
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
O=gpurun_out/t_ab.txt; : > $O
for g in 1 0; do
  echo "== grid=$g headline" >> $O; RTB_GRID=$g timeout 300 python tools/prof_frame.py --reps 3 --count >> $O 2>&1
  echo "== grid=$g 10k" >> $O; RTB_GRID=$g timeout 300 python tools/prof_frame.py --workload spheres10k_1080p_64spp_d16 --reps 3 --count >> $O 2>&1
done
CMD="python tools/prof_frame.py --spp 32 --reps 1"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:render_mega -c 1 -f -o /tmp/megag $CMD > gpurun_out/t_ncu.log 2>&1; echo "ncu rc=$?"
python tools/ncu_summary.py /tmp/megag.ncu-rep > gpurun_out/t_ncu_grid_summary.txt 2>&1
python tools/ncu_by_line.py /tmp/megag.ncu-rep ray-tracing-practice_b200/librt_b200.so render_megaILi0ELb0ELb0ELb0ELb1E --top 400 > gpurun_out/t_ncu_grid_by_line.txt 2>&1
grep -E "^==|kernel_ms" $O | sed 's/{"workload".*"kernel_ms"/ms/; s/"mean".*"per_ray"/per_ray/' | paste - - | cut -c1-330
head -14 gpurun_out/t_ncu_grid_by_line.txt
