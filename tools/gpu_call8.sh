#!/bin/bash
# r01 verification call: GPU tests, smoke, per-kernel bench, full bench (both arms), ncu launch list + --set full captures.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 600 python tools/kbench.py > gpurun_out/kbench.log 2>&1; echo "kbench rc=$?"
cat gpurun_out/kbench.log | cut -c1-200
timeout 1200 python bench.py --warmup 1 > gpurun_out/bench_w1.json 2> gpurun_out/bench_w1.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/bench_w1.json; tail -3 gpurun_out/bench_w1.err
timeout 300 python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
# launch list: full-size tensors, 5 SPLOCS iterations (the 50-iteration command has 16.5k launches: too slow under ncu)
LARGS="--steps 1 --warmup 0 --no-e2e --no-cpu-baseline --splocs-iters 5"
timeout 600 python bench.py $LARGS > gpurun_out/bench_is5_plain.json 2> gpurun_out/bench_is5_plain.err &&
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 30000 --csv --log-file gpurun_out/launches_is5.csv \
    python bench.py $LARGS > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/launches_is5.csv
# --set full: one launch of each hot kernel at the full 1M x 2000 shape
timeout 300 python tools/kbench.py --once --only stream,gemm,admm > gpurun_out/kbench_once.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on \
    -k regex:"deflate_gemv_kernel|rank_update_kernel|gemm_nt_skinny|gemm_nn_skinny|admm_solve_kernel|gemm_nt_kernel|gemm_nn_kernel" -c 8 \
    -o gpurun_out/prof_hot -f python tools/kbench.py --once --only stream,gemm,admm > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -3 gpurun_out/ncu_full.log; ls -la gpurun_out
