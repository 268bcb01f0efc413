#!/bin/bash
# ADMM layout sweep (numerics + speed), then the full bench with the fastest
set -u
mkdir -p gpurun_out
for L in 4 416 12 1216; do
  ASB_ADMM_LAYOUT=$L timeout 600 python -m pytest tests -m gpu -q -k "admm or splocs or pipeline or protocol" > gpurun_out/pytest_L$L.log 2>&1; echo "pytest layout $L rc=$?"; tail -1 gpurun_out/pytest_L$L.log
done
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest default rc=$?"; tail -1 gpurun_out/pytest_gpu.log
for CFG in "ASB_ADMM_LAYOUT=4 ASB_ADMM_LOCK=1" "ASB_ADMM_LAYOUT=4 ASB_ADMM_LOCK=0" "ASB_ADMM_LAYOUT=416 ASB_ADMM_LOCK=1" "ASB_ADMM_LAYOUT=416 ASB_ADMM_LOCK=0" "ASB_ADMM_LAYOUT=12" "ASB_ADMM_LAYOUT=1216"; do
  TAG=$(echo $CFG | tr -d ' ' | tr '=' '_')
  env $CFG timeout 300 python tools/kbench.py --only admm > gpurun_out/kb_$TAG.log 2>&1
  echo "$CFG: $(grep admm_solve_x10 gpurun_out/kb_$TAG.log | cut -c1-120)"
done
BEST=$(python - <<'PY'
import json, glob
best=None
for f in glob.glob("gpurun_out/kb_ASB_ADMM_LAYOUT*.log"):
    try:
        ms=[json.loads(l)["ms"] for l in open(f) if "admm_solve_x10" in l][-1]
    except Exception:
        continue
    tag=f.split("kb_")[1][:-4]
    env=tag.replace("ASB_ADMM_LAYOUT_","ASB_ADMM_LAYOUT=").replace("ASB_ADMM_LOCK_"," ASB_ADMM_LOCK=")
    if best is None or ms<best[0]: best=(ms,env)
print(best[1] if best else "ASB_ADMM_LAYOUT=4")
PY
)
echo "best ADMM config: $BEST"
env $BEST timeout 900 python bench.py --warmup 0 --steps 1 --no-e2e --no-cpu-baseline > gpurun_out/bench_call11.json 2> gpurun_out/bench_call11.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_call11.json').read().strip().splitlines()[-1])
print(d['value'], d['roofline']['achieved'], d['path_roofline']['frac']); print(d['kernel_ms_per_step'])
PY
tail -3 gpurun_out/bench_call11.err
