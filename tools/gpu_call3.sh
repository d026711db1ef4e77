#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/c3_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/c3_pytest.log
V=$GRAFT_REPO_ROOT/ray-tracing-practice_b200/build
echo "== r01 headline" >> gpurun_out/c3_ab.txt; (cd build/r01 && timeout 300 python tools/prof_frame.py --reps 3) >> gpurun_out/c3_ab.txt 2>&1
echo "== r01 shipped" >> gpurun_out/c3_ab.txt; (cd build/r01 && timeout 300 python tools/prof_frame.py --workload shipped --reps 2) >> gpurun_out/c3_ab.txt 2>&1
for v in default v_planeexact v_fastD; do
  if [ $v = default ]; then unset RTB_LIB; else export RTB_LIB=$V/$v.so; fi
  echo "== $v headline" >> gpurun_out/c3_ab.txt; timeout 300 python tools/prof_frame.py --reps 3 >> gpurun_out/c3_ab.txt 2>&1
  echo "== $v shipped" >> gpurun_out/c3_ab.txt; timeout 300 python tools/prof_frame.py --workload shipped --reps 2 >> gpurun_out/c3_ab.txt 2>&1
done
unset RTB_LIB
echo "== default headline exact" >> gpurun_out/c3_ab.txt; timeout 300 python tools/prof_frame.py --reps 2 --exact >> gpurun_out/c3_ab.txt 2>&1
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/c3_bench.json 2> gpurun_out/c3_bench.err; echo "bench rc=$?"
grep -E "passed|failed|rc=" gpurun_out/c3_pytest.log | tail -5
grep -E "^==|kernel_ms" gpurun_out/c3_ab.txt | sed 's/"segments.*//' | cut -c1-160
