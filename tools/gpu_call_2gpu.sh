#!/bin/bash
# 2-GPU check: sharded golden pipeline over NCCL, then the row-sharded bench at the full workload
set -u
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29621 \
    tools/nccl_worker.py > gpurun_out/nccl_worker.log 2>&1; echo "nccl worker rc=$?"; grep "OK\|Error\|error" gpurun_out/nccl_worker.log | tail -5
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29622 \
    bench.py --gpus 2 --warmup 0 --steps 1 --no-cpu-baseline > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; echo "bench2 rc=$?"
tail -c 2500 gpurun_out/bench_2gpu.json; tail -5 gpurun_out/bench_2gpu.err
