#!/bin/bash
# 8-GPU box: full GPU tier (multi-GPU cases included), bench at N = 8 / 4 / 2 / 1, the 8K config at N = 8, f3 frames over 8 GPUs
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/p_gpus.txt
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/p_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/p_pytest.log
for n in 8 4 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/p_bench_n$n.json 2> gpurun_out/p_bench_n$n.err; echo "bench$n rc=$?"
done
timeout 600 python bench.py --steps 5 --warmup 3 --no-baselines > gpurun_out/p_bench_n1.json 2> gpurun_out/p_bench_n1.err; echo "bench1 rc=$?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus 8 --steps 3 --warmup 3 --workload spheres100k_8k_1024spp_d32 > gpurun_out/p_bench_n8_8k.json 2> gpurun_out/p_bench_n8_8k.err; echo "bench 8k rc=$?"
timeout 900 python tools/time_frames.py --frames 100 --gpus 8 --only pipelined_frames_over,pipelined_tiles_over > gpurun_out/p_frames_8gpu.jsonl 2> gpurun_out/p_frames_8gpu.err; echo "frames rc=$?"
tail -3 gpurun_out/p_pytest.log
for f in gpurun_out/p_bench_n8.json gpurun_out/p_bench_n4.json gpurun_out/p_bench_n2.json gpurun_out/p_bench_n1.json gpurun_out/p_bench_n8_8k.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); print(sys.argv[1], round(d['value']), 'Mrays/s', round(d['ms_per_step'],2),'ms e2e', round(d['e2e']['value']), d.get('per_rank',{}).get('kernel_ms'))
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
done
cut -c1-330 gpurun_out/p_frames_8gpu.jsonl
