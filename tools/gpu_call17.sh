#!/bin/bash
# ADMM Tall layout (16 warps: 8 row-warps x 2 vertex groups, U through L2) vs the default Pair
set -u
mkdir -p gpurun_out
ASB_ADMM_LAYOUT=16 timeout 600 python -m pytest tests -m gpu -q -k "admm or splocs or pipeline or protocol" > gpurun_out/pytest_L16.log 2>&1; echo "pytest layout 16 rc=$?"; tail -1 gpurun_out/pytest_L16.log
for L in 16 4; do
  ASB_ADMM_LAYOUT=$L timeout 300 python tools/kbench.py --only admm > gpurun_out/kb_L$L.log 2>&1
  echo "layout $L: $(grep admm_solve_x10 gpurun_out/kb_L$L.log | cut -c1-120)"
done
ASB_ADMM_LAYOUT=16 timeout 900 python bench.py --warmup 0 --steps 1 --no-e2e --no-cpu-baseline > gpurun_out/bench_L16.json 2> gpurun_out/bench_L16.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench_L16.json') if l.startswith('{')][-1])
print(d['value'], d['roofline']['achieved'], d['path_roofline']['frac'], d['phase_ms_per_step']); print(d['kernel_ms_per_step'])
PY
