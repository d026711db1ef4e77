#!/bin/bash
# what the driver runs at round end, on one GPU: GPU tier, smoke(), reference arm, bench
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/z_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/z_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/z_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/z_smoke.log
timeout 600 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 > gpurun_out/z_bench_reference_arm.json 2> gpurun_out/z_bench_reference_arm.err; echo "ref arm rc=$?"
timeout 600 python bench.py > gpurun_out/z_bench.json 2> gpurun_out/z_bench.err; echo "bench rc=$?"
wc -l gpurun_out/z_bench.json gpurun_out/z_bench_reference_arm.json
tail -2 gpurun_out/z_pytest.log; tail -2 gpurun_out/z_smoke.log; cut -c1-200 gpurun_out/z_bench.json
