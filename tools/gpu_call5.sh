#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/pytest_gpu.log
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_ops.py -x -q -k "admm or gemm" > gpurun_out/memcheck.log 2>&1; echo "memcheck rc=$?"
grep -E "ERROR SUMMARY|passed|failed" gpurun_out/memcheck.log | tail -3
timeout 900 python bench.py --warmup 1 > gpurun_out/bench_w1.json 2> gpurun_out/bench_w1.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_w1.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','e2e','roofline','path_roofline','kernel_ms_per_step','gpu_launches')})
PY
# launch list on a reduced workload (same command run plainly first)
RED="--steps 1 --warmup 0 --no-e2e --no-cpu-baseline --grid 300 300 --frames 500 --comps 48 --splocs-iters 3"
timeout 300 python bench.py $RED > gpurun_out/bench_reduced.json 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file gpurun_out/launches_reduced.csv \
    python bench.py $RED > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/launches_reduced.csv
# full capture of the fused ADMM kernel on a shard-sized problem
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"admm_solve" -c 2 \
    -o gpurun_out/prof_admm -f python tools/kbench.py --verts 200000 --only admm > gpurun_out/ncu_admm.log 2>&1
echo "ncu admm rc=$?"; ls -la gpurun_out/*.ncu-rep
