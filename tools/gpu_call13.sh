#!/bin/bash
# r01 closing measurement: GPU suite, smoke, default bench (both arms), dram traffic of the dominant kernel, launch list
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 300 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 1500 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_default.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e'], d['roofline']['achieved'], d['path_roofline']['frac'], d['phase_ms_per_step']); print(d['kernel_ms_per_step'])
PY
tail -3 gpurun_out/bench_default.err
# dram traffic of the dominant kernel at the full shape (3 metrics only: one or two replay passes)
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
    -k regex:"admm_solve_kernel|gemm_nt_skinny|gemm_nn_skinny" -c 4 --csv --log-file gpurun_out/traffic_hot.csv \
    python tools/kbench.py --once --only gemm,admm > gpurun_out/ncu_traffic.log 2>&1; echo "traffic rc=$?"
grep -v "^==" gpurun_out/traffic_hot.csv | cut -d, -f5,13- | head -14
# launch list of a bounded command on the full tensor (48 components, 3 SPLOCS iterations), after its plain run
LARGS="--steps 1 --warmup 0 --no-e2e --no-cpu-baseline --comps 48 --splocs-iters 3"
timeout 600 python bench.py $LARGS > gpurun_out/bench_k48_plain.json 2> gpurun_out/bench_k48_plain.err &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:"asb::" -c 30000 --csv \
    --log-file gpurun_out/launches_k48.csv python bench.py $LARGS > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/launches_k48.csv
