#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
CMD="python tools/prof_frame.py --reps 1"
$CMD > gpurun_out/c8_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:render_mega -c 1 -f -o gpurun_out/c8_prof $CMD > gpurun_out/c8_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/c8_ncu.log; ls -la gpurun_out/c8_prof.ncu-rep
