#!/bin/bash
# the last ~20 s of the r01 GPU budget: the snapshot-producer kernel and the zero-residual check
timeout 9 python -m pytest tests/test_zz_gpu_late.py -x -q -rxX -p no:cacheprovider -k "pd_local or zero_residual" > gpurun_out/zz_late2.log 2>&1
echo "rc=$?" >> gpurun_out/zz_late2.log
tail -6 gpurun_out/zz_late2.log
