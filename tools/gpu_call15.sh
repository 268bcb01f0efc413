#!/bin/bash
# geodesic factorisation leaf size: 32 and 16 against the default 64 (17.6 s)
set -u
mkdir -p gpurun_out
for L in 32 16; do
  timeout 600 python bench.py --warmup 0 --steps 1 --no-e2e --no-cpu-baseline --leaf-size $L > gpurun_out/bench_leaf$L.json 2> gpurun_out/bench_leaf$L.err; echo "leaf $L rc=$?"
  python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/bench_leaf$L.json') if l.startswith('{')][-1])
print($L, d['value'], d['phase_ms_per_step'], d['geodesic'], d['kernel_ms_per_step']['asb_mf_solve'])
PY
done
