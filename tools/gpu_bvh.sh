#!/bin/bash
# device BVH build: GPU tests, memcheck/racecheck of the small cases, headline frame with both builders (+ work counters)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_bvh_build.py -x -q -s > gpurun_out/pytest_bvh.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_bvh.log
for b in host gpu; do
  timeout 300 python tools/prof_frame.py --builder $b --count --reps 3 > gpurun_out/frame_$b.json 2>&1
done
timeout 300 python tools/prof_frame.py --workload spheres1m_4k_64spp_d8 --builder gpu --spp 8 --count --reps 2 > gpurun_out/frame_1m_gpu.json 2>&1
timeout 300 python tools/prof_frame.py --workload spheres1m_4k_64spp_d8 --builder host --spp 8 --count --reps 2 > gpurun_out/frame_1m_host.json 2>&1
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
