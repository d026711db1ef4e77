#!/bin/bash
# final records: default bench line, reference arm, every single-GPU config through bench.py (reference CUDA kernel beside it), ncu of the final kernel
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/q_bench_final.json 2> gpurun_out/q_bench_final.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/q_bench_reference_arm.json 2> gpurun_out/q_bench_reference_arm.err; echo "ref arm rc=$?"
: > gpurun_out/q_all_configs_bench.jsonl
for w in test_config_200x100 spheres10k_1080p_64spp_d16 spheres1m_4k_64spp_d8 shipped_textured_1080x720_2500spp_d50; do
  timeout 900 python bench.py --workload $w --steps 3 --warmup 3 >> gpurun_out/q_all_configs_bench.jsonl 2>> gpurun_out/q_all_configs_bench.err; echo "$w rc=$?"
done
CMD="python tools/prof_frame.py --reps 1"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:render_mega -c 1 -f -o /tmp/mega $CMD > gpurun_out/q_ncu.log 2>&1; echo "ncu rc=$?"
python tools/ncu_summary.py /tmp/mega.ncu-rep --json gpurun_out/q_ncu_mega_headline.json > gpurun_out/q_ncu_mega_summary.txt 2>&1
python tools/ncu_by_line.py /tmp/mega.ncu-rep ray-tracing-practice_b200/librt_b200.so render_megaILi0ELb0ELb0ELb0E --top 120 > gpurun_out/q_ncu_mega_by_source_line.txt 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/q_launches.csv python bench.py --steps 2 --warmup 1 --no-baselines > gpurun_out/q_launches.log 2>&1
cut -c1-300 gpurun_out/q_bench_final.json; cut -c1-200 gpurun_out/q_all_configs_bench.jsonl
