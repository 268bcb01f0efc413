#!/bin/bash
# ADMM: B0 re-loaded inside the epilogue (no exposed L2 latency in front of the first DMMA of an iteration)
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -k "admm or splocs or pipeline or protocol" > gpurun_out/pytest_admm.log 2>&1; echo "pytest rc=$?"; tail -1 gpurun_out/pytest_admm.log
ASB_ADMM_LAYOUT=16 timeout 600 python -m pytest tests -m gpu -q -k "admm" > gpurun_out/pytest_admm16.log 2>&1; echo "pytest16 rc=$?"; tail -1 gpurun_out/pytest_admm16.log
for L in 4 16 12; do
  ASB_ADMM_LAYOUT=$L timeout 300 python tools/kbench.py --only admm > gpurun_out/kb_L$L.log 2>&1
  echo "layout $L: $(grep admm_solve_x10 gpurun_out/kb_L$L.log | cut -c1-120)"
done
