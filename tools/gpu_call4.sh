#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -8 gpurun_out/pytest_gpu.log
timeout 600 python tools/kbench.py --only gemm,admm > gpurun_out/kbench.log 2>&1; echo "kbench rc=$?"
tail -8 gpurun_out/kbench.log
