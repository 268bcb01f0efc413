#!/bin/bash
# ADMM Duo layouts: 2 CTAs/SM x 8 row-warps (16 warps per SM in two independent CTAs)
mkdir -p gpurun_out
ASB_ADMM_LAYOUT=8 timeout 60 python -m pytest tests/test_gpu_ops.py -m gpu -q -k "admm" 2>&1 | tail -1
for L in 8 81; do
  ASB_ADMM_LAYOUT=$L timeout 60 python tools/kbench.py --only admm > gpurun_out/kb_L$L.log 2>&1
  echo "layout $L: $(grep admm_solve_x10 gpurun_out/kb_L$L.log | cut -c1-110)"
done
