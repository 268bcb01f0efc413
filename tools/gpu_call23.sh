#!/bin/bash
# ncu --set full of the ADMM kernel, layouts 4 (Pair) and 16 (Tall), 200k vertices (cheap in-place restore)
set -u
mkdir -p gpurun_out
for L in 4 16; do
  ASB_ADMM_LAYOUT=$L timeout 200 ncu --set full --clock-control none --import-source on -k regex:"admm_solve_kernel" -c 1 \
      -o gpurun_out/prof_admm_L$L -f python tools/kbench.py --verts 200000 --once --only admm > gpurun_out/ncu_admm_L$L.log 2>&1
  echo "ncu layout $L rc=$?"
done
ls -la gpurun_out/*.ncu-rep
