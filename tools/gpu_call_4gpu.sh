#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29624 \
    bench.py --gpus 4 --warmup 1 --steps 1 --no-e2e --no-cpu-baseline > gpurun_out/bench_4gpu.json 2> gpurun_out/bench_4gpu.err; echo "bench4 rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_4gpu.json') if l.startswith('{')][-1])
print(d['value'], d['phase_ms_per_step'], d['geodesic']); print(d['kernel_ms_per_step'])
PY
tail -3 gpurun_out/bench_4gpu.err
