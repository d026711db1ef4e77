#!/bin/bash
# ncu --set full of render_mega (headline frame, one launch) + launch list of the bench command
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
CMD="python tools/prof_frame.py --reps 1"
$CMD > gpurun_out/m_plain.log 2>&1 || exit 1
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:render_mega -c 1 -f -o /tmp/mega $CMD > gpurun_out/m_ncu.log 2>&1; echo "ncu rc=$?"
python tools/ncu_summary.py /tmp/mega.ncu-rep --json gpurun_out/m_ncu_mega_headline.json > gpurun_out/m_ncu_mega_summary.txt 2>&1
python tools/ncu_by_line.py /tmp/mega.ncu-rep ray-tracing-practice_b200/librt_b200.so render_megaILi0ELb0ELb0ELb0E --top 500 > gpurun_out/m_ncu_mega_by_source_line.txt 2>&1
CMD2="python tools/prof_frame.py --workload shipped --spp 256 --reps 1"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_mega -c 1 -f -o /tmp/megas $CMD2 > gpurun_out/m_ncu_s.log 2>&1
python tools/ncu_summary.py /tmp/megas.ncu-rep > gpurun_out/m_ncu_mega_shipped_summary.txt 2>&1
python tools/ncu_by_line.py /tmp/megas.ncu-rep ray-tracing-practice_b200/librt_b200.so render_megaILi0ELb0ELb0ELb0E --top 500 > gpurun_out/m_ncu_mega_shipped_by_source_line.txt 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/m_launches.csv python bench.py --steps 2 --warmup 1 --no-baselines > gpurun_out/m_launches.log 2>&1
head -50 gpurun_out/m_ncu_mega_summary.txt
