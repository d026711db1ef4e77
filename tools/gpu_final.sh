#!/bin/bash
# final measurement of the round on one GPU: suite, smoke, bench (both arms), all configs, launch list, full ncu capture
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?" >> gpurun_out/bench_n1.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 900 python tools/run_configs.py > gpurun_out/all_configs.jsonl 2> gpurun_out/all_configs.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-baselines > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:render_mega -c 1 -o gpurun_out/prof_mega_head python tools/prof_frame.py --reps 1 > gpurun_out/ncu_full.log 2>&1
timeout 300 python tools/prof_frame.py --builder gpu --reps 2 --count > gpurun_out/frame_gpu_builder.json 2>&1
timeout 300 python tools/prof_frame.py --workload spheres1m_4k_64spp_d8 --builder gpu --reps 2 --count > gpurun_out/frame_1m_gpu_builder.json 2>&1
ls -la gpurun_out
