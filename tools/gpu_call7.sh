#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -x -q -k "gemm" > gpurun_out/pytest_wn2.log 2>&1; echo "pytest wn2 rc=$?"
ASB_SK_WN=3 timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -x -q -k "gemm" > gpurun_out/pytest_wn3.log 2>&1; echo "pytest wn3 rc=$?"
tail -3 gpurun_out/pytest_wn3.log
ASB_SK_WN=2 timeout 300 python tools/kbench.py --only gemm > gpurun_out/kbench_wn2.log 2>&1; tail -3 gpurun_out/kbench_wn2.log
ASB_SK_WN=3 timeout 300 python tools/kbench.py --only gemm > gpurun_out/kbench_wn3.log 2>&1; tail -3 gpurun_out/kbench_wn3.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"admm_solve" --launch-skip 5 -c 1 \
    -o gpurun_out/prof_admm10 -f python tools/kbench.py --verts 200000 --only admm > gpurun_out/ncu_admm.log 2>&1
echo "ncu admm rc=$?"; ls -la gpurun_out/prof_admm10.ncu-rep
