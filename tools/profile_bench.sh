#!/bin/bash
# Run under gpurun on ONE B200:  plain bench (must exit 0), then the ncu launch list of the same command, then one
# `--set full` capture of the dominant kernel.  Outputs land in gpurun_out/ (copy summaries into profiles/).
set -u
ARGS="--steps 1 --warmup 0 --no-e2e --no-cpu-baseline $*"
mkdir -p gpurun_out
python bench.py $ARGS > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 40000 --csv --log-file gpurun_out/launches.csv \
    python bench.py $ARGS > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
tail -2 gpurun_out/plain.log
python bench.py $ARGS > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:${KERNEL:-admm_iter} -s ${SKIP:-3} -c 2 --kill yes \
    -o gpurun_out/prof_${KERNEL:-admm_iter} -f python bench.py $ARGS > gpurun_out/ncu_full.log 2>&1
echo "full capture rc=$?"
tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out
