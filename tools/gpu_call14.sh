#!/bin/bash
# 2 GPUs: sharded golden pipeline over NCCL (geodesic all-to-all on the side stream), sharded bench; rank-0-only 1-GPU bench for the priority stream
set -u
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29621 \
    tools/nccl_worker.py > gpurun_out/nccl_worker.log 2>&1; echo "nccl worker rc=$?"; grep "OK\|Error\|error" gpurun_out/nccl_worker.log | tail -5
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29622 \
    bench.py --gpus 2 --warmup 1 --steps 1 --no-cpu-baseline > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; echo "bench2 rc=$?"
CUDA_VISIBLE_DEVICES=0 timeout 600 python bench.py --warmup 0 --steps 1 --no-e2e --no-cpu-baseline > gpurun_out/bench_prio.json 2> gpurun_out/bench_prio.err; echo "bench1 rc=$?"
python - <<'PY'
import json
for f in ('gpurun_out/bench_2gpu.json','gpurun_out/bench_prio.json'):
    d=json.loads([l for l in open(f) if l.startswith('{')][-1])
    print(f, d['value'], d['e2e'], d['phase_ms_per_step'], d['geodesic']); print(d['kernel_ms_per_step'])
PY
tail -3 gpurun_out/bench_2gpu.err
