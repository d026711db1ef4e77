#!/bin/bash
# state of HEAD: full GPU test tier, default bench line, reference arm
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x -s > gpurun_out/l_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/l_pytest.log
timeout 600 python bench.py > gpurun_out/l_bench.json 2> gpurun_out/l_bench.err; echo "bench rc=$?"
tail -5 gpurun_out/l_pytest.log
cut -c1-600 gpurun_out/l_bench.json
