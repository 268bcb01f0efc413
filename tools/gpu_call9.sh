#!/bin/bash
# multi-RHS multifrontal solve check + ADMM ncu capture (200k vertices: in-place Z/U restore is cheap) + full bench
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops.py tests/test_gpu_pipeline.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python tools/kbench.py --verts 200000 --once --only admm > gpurun_out/kbench_admm200k.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"admm_solve_kernel" -c 1 \
    -o gpurun_out/prof_admm -f python tools/kbench.py --verts 200000 --once --only admm > gpurun_out/ncu_admm.log 2>&1
echo "ncu admm rc=$?"; tail -2 gpurun_out/ncu_admm.log
timeout 900 python bench.py --warmup 0 --steps 1 --no-e2e --no-cpu-baseline > gpurun_out/bench_mrhs.json 2> gpurun_out/bench_mrhs.err; echo "bench rc=$?"
tail -c 1800 gpurun_out/bench_mrhs.json; tail -3 gpurun_out/bench_mrhs.err
