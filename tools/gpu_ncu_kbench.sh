#!/bin/bash
# ncu --set full of the DMMA kernels on a 200k-vertex shard-sized problem (kbench), one launch each.
set -u
mkdir -p gpurun_out
ARGS="tools/kbench.py --verts ${VERTS:-200000} --only ${ONLY:-gemm,admm}"
timeout 300 python $ARGS > gpurun_out/kbench_small.log 2>&1 || { echo "plain kbench failed"; tail -5 gpurun_out/kbench_small.log; exit 1; }
cat gpurun_out/kbench_small.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"${KREGEX:-skinny|admm_solve}" -c ${COUNT:-12} \
    -o gpurun_out/prof_dmma -f python $ARGS > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_full.log; ls -la gpurun_out/*.ncu-rep
