#!/bin/bash
# full GPU tier (new tests: grid scene, schedule invariance, signed bias) + bench line
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x -s > gpurun_out/u_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/u_pytest.log
timeout 600 python bench.py --workload shipped_textured_1080x720_2500spp_d50 --steps 3 --warmup 3 > gpurun_out/u_bench_shipped.json 2> gpurun_out/u_bench_shipped.err; echo "bench shipped rc=$?"
tail -4 gpurun_out/u_pytest.log | cut -c1-300; grep -n "signed bias" gpurun_out/u_pytest.log
cut -c1-250 gpurun_out/u_bench_shipped.json
