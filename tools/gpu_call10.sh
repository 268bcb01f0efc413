#!/bin/bash
# ADMM layout A/B (Wide 12-warp vs Pair 2x4-warp with/without the epilogue mutex) + DMMA forward multifrontal solve, then the full bench
set -u
mkdir -p gpurun_out
K="admm or splocs or pipeline or multifrontal or geodes or smoke"
ASB_ADMM_LAYOUT=12 timeout 600 python -m pytest tests -m gpu -x -q -k "$K" > gpurun_out/pytest_wide.log 2>&1; echo "pytest wide rc=$?"; tail -2 gpurun_out/pytest_wide.log
ASB_ADMM_LAYOUT=4 ASB_ADMM_LOCK=1 timeout 600 python -m pytest tests -m gpu -x -q -k "admm or splocs or pipeline" > gpurun_out/pytest_pair.log 2>&1; echo "pytest pair rc=$?"; tail -2 gpurun_out/pytest_pair.log
ASB_ADMM_LAYOUT=12 timeout 300 python tools/kbench.py --only admm > gpurun_out/kb_wide.log 2>&1; tail -1 gpurun_out/kb_wide.log | cut -c1-160
ASB_ADMM_LAYOUT=4 ASB_ADMM_LOCK=1 timeout 300 python tools/kbench.py --only admm > gpurun_out/kb_pair_lock.log 2>&1; tail -1 gpurun_out/kb_pair_lock.log | cut -c1-160
ASB_ADMM_LAYOUT=4 ASB_ADMM_LOCK=0 timeout 300 python tools/kbench.py --only admm > gpurun_out/kb_pair_nolock.log 2>&1; tail -1 gpurun_out/kb_pair_nolock.log | cut -c1-160
BEST=$(python - <<'PY'
import json
best=None
for name,env in (("kb_wide","ASB_ADMM_LAYOUT=12"),("kb_pair_lock","ASB_ADMM_LAYOUT=4 ASB_ADMM_LOCK=1"),("kb_pair_nolock","ASB_ADMM_LAYOUT=4 ASB_ADMM_LOCK=0")):
    try:
        ms=[json.loads(l)["ms"] for l in open(f"gpurun_out/{name}.log") if "admm_solve_x10" in l][-1]
    except Exception:
        continue
    if best is None or ms<best[0]: best=(ms,env)
print(best[1] if best else "ASB_ADMM_LAYOUT=4 ASB_ADMM_LOCK=0")
PY
)
echo "best ADMM config: $BEST"
env $BEST timeout 900 python bench.py --warmup 0 --steps 1 --no-e2e --no-cpu-baseline > gpurun_out/bench_call10.json 2> gpurun_out/bench_call10.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_call10.json').read().strip().splitlines()[-1])
print(d['value'], d['roofline'], d['path_roofline']['frac']); print(d['kernel_ms_per_step'])
PY
tail -3 gpurun_out/bench_call10.err
