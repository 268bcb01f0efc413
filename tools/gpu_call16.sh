#!/bin/bash
# geometric DEIM + DEIM reconstruction on the device: full GPU suite (1 GPU) and the NCCL-sharded golden checks (2 GPUs)
set -u
mkdir -p gpurun_out
CUDA_VISIBLE_DEVICES=0 timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29621 \
    tools/nccl_worker.py > gpurun_out/nccl_worker.log 2>&1; echo "nccl worker rc=$?"; grep "OK\|Error\|error\|assert" gpurun_out/nccl_worker.log | tail -8
