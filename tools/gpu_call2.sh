#!/bin/bash
# parity tests, per-kernel timings of the DMMA kernels, one full bench (device-resident + e2e + cpu baseline)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python tools/kbench.py --only gemm,admm,dgemm > gpurun_out/kbench.log 2>&1; echo "kbench rc=$?"
cat gpurun_out/kbench.log | tail -12
timeout 900 python bench.py --warmup 1 > gpurun_out/bench_w1.json 2> gpurun_out/bench_w1.err; echo "bench rc=$?"
tail -c 2500 gpurun_out/bench_w1.json; tail -5 gpurun_out/bench_w1.err
