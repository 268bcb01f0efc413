#!/bin/bash
# One parameterised GPU session (gpurun -- bash tools/gpu_session.sh <steps...>): every step writes its log under gpurun_out/.
# steps: tests | scale | geo | bench1 | bench <args> | kbench <args>
mkdir -p gpurun_out
for step in "$@"; do
  case "$step" in
    tests)  timeout 1500 python -m pytest tests -x -q -m gpu -rxXs > gpurun_out/tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/tests.log ;;
    scale)  timeout 1500 python -m pytest tests/test_gpu_scale.py -q -m gpu -rxXs > gpurun_out/scale.log 2>&1; echo "scale rc=$?"; tail -15 gpurun_out/scale.log ;;
    geo)    for t in 1 0; do ASB_MF_TALL=$t timeout 600 python tools/kbench.py --geo 1000 1000 >> gpurun_out/geo.log 2>&1; done; cat gpurun_out/geo.log | tail -4 ;;
    bench1) timeout 900 python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/bench1.json 2> gpurun_out/bench1.err; echo "bench1 rc=$?"; tail -c 3000 gpurun_out/bench1.json; tail -5 gpurun_out/bench1.err ;;
    *) echo "unknown step $step" ;;
  esac
done
