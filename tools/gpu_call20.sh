#!/bin/bash
# delta-source heat solves along the tree path: parity suite + bench
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --warmup 0 --steps 1 --no-e2e --no-cpu-baseline > gpurun_out/bench_call20.json 2> gpurun_out/bench_call20.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_call20.json') if l.startswith('{')][-1])
print(d['value'], d['path_roofline']['frac'], d['phase_ms_per_step']); print(d['kernel_ms_per_step'])
PY
tail -3 gpurun_out/bench_call20.err
