#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/c2_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/c2_pytest.log
for c in 1 2 4 8 32; do for m in 0 1; do
  echo "== headline chunk=$c major=$m" >> gpurun_out/c2_chunks.txt
  RTB_SAMPLE_CHUNK=$c RTB_CHUNK_MAJOR=$m timeout 300 python tools/prof_frame.py --reps 3 >> gpurun_out/c2_chunks.txt 2>&1
done; done
for c in 1 2 4 20; do
  echo "== shipped chunk=$c" >> gpurun_out/c2_chunks.txt
  RTB_SAMPLE_CHUNK=$c timeout 300 python tools/prof_frame.py --workload shipped --reps 2 >> gpurun_out/c2_chunks.txt 2>&1
done
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/c2_bench.json 2> gpurun_out/c2_bench.err; echo "bench rc=$?"
grep -E "passed|failed|rc=" gpurun_out/c2_pytest.log | tail -5
grep -E "^==|kernel_ms" gpurun_out/c2_chunks.txt | sed 's/"segments.*//' | cut -c1-160
