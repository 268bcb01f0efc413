#!/bin/bash
# last seconds of the r01 GPU budget: the late GPU tests that do not run the oracle live
timeout 33 python -m pytest tests/test_zz_gpu_late.py -x -q -k "not live_oracle" > gpurun_out/zz_late.log 2>&1
echo "rc=$?" >> gpurun_out/zz_late.log
tail -5 gpurun_out/zz_late.log
