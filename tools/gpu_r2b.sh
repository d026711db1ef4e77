#!/bin/bash
# no-alloc entry point: alone in a fresh process, then after the whole GPU tier
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/b_noalloc_alone.log 2>&1; echo "alone rc=$?"
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/b_pytest.log 2>&1; echo "all rc=$?"
tail -15 gpurun_out/b_noalloc_alone.log; tail -4 gpurun_out/b_pytest.log
