#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/pytest_gpu.log
timeout 600 python tools/kbench.py --only admm > gpurun_out/kbench.log 2>&1; echo "kbench rc=$?"
tail -3 gpurun_out/kbench.log
timeout 900 python bench.py --warmup 1 > gpurun_out/bench_w1.json 2> gpurun_out/bench_w1.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_w1.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','e2e','roofline','path_roofline','kernel_ms_per_step','kernel_launches_per_step','gpu_launches','geodesic')})
PY
tail -3 gpurun_out/bench_w1.err
