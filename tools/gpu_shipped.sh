#!/bin/bash
# the reference's own shipped scene: work counters and a full ncu capture of a 256-spp frame
mkdir -p gpurun_out
timeout 300 python tools/prof_frame.py --workload shipped --spp 256 --count --reps 2 > gpurun_out/shipped_frame.json 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_mega -c 1 -o gpurun_out/prof_mega_shipped python tools/prof_frame.py --workload shipped --spp 256 --reps 1 > gpurun_out/ncu_shipped.log 2>&1
