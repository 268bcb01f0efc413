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
    bench1e) timeout 1200 python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/bench1e.json 2> gpurun_out/bench1e.err; echo "bench1e rc=$?"; tail -c 3500 gpurun_out/bench1e.json; tail -5 gpurun_out/bench1e.err ;;
    presets) for c in bunny_splocs armadillo_splocs bar_deim synthetic_sample; do timeout 600 python bench.py --config $c --steps 3 --warmup 3 > gpurun_out/bench_$c.json 2> gpurun_out/bench_$c.err; echo "$c rc=$?"; tail -c 1800 gpurun_out/bench_$c.json; tail -3 gpurun_out/bench_$c.err; done ;;
    u2)     timeout 900 python bench.py --unit pod_deim --steps 2 > gpurun_out/bench_pod_deim.json 2> gpurun_out/bench_pod_deim.err; echo "u2 rc=$?"; tail -c 3000 gpurun_out/bench_pod_deim.json; tail -5 gpurun_out/bench_pod_deim.err ;;
    trace)  for pr in 0 -1; do echo "== side priority $pr"; ASB_SIDE_PRIORITY=$pr ASB_TRACE_DEFLATE=1 timeout 600 python bench.py --splocs-iters 0 --steps 2 --warmup 1 --no-e2e --no-cpu-baseline 2>&1 | grep -o "\[deflate trace.*\|\"phase_ms_per_step\": {[^}]*}" | tail -3; done ;;
    admm)   if ASB_ADMM_LAYOUT=5 timeout 180 python -m pytest tests/test_gpu_ops.py -x -q -m gpu -k admm > gpurun_out/admm5_tests.log 2>&1; then
              tail -2 gpurun_out/admm5_tests.log
              for l in 5 4; do ASB_ADMM_LAYOUT=$l timeout 300 python tools/kbench.py --only admm 2>&1 | grep admm_solve_x10 | sed "s/^/layout $l: /"; done
            else echo "layout 5 tests FAILED"; tail -30 gpurun_out/admm5_tests.log; fi ;;
    duet)   for sk in 1 2 0; do
              if ASB_ADMM_LAYOUT=6 ASB_ADMM_SKEW=$sk timeout 180 python -m pytest tests/test_gpu_ops.py -x -q -m gpu -k admm > gpurun_out/admm6_$sk.log 2>&1; then
                ASB_ADMM_LAYOUT=6 ASB_ADMM_SKEW=$sk timeout 300 python tools/kbench.py --only admm 2>&1 | grep admm_solve_x10 | sed "s/^/duet skew $sk: /"
              else echo "duet skew $sk tests FAILED"; tail -20 gpurun_out/admm6_$sk.log; fi
            done
            ASB_ADMM_LAYOUT=5 ASB_ADMM_TMA_SOLO=1 timeout 300 python tools/kbench.py --only admm 2>&1 | grep admm_solve_x10 | sed "s/^/tma solo (1 CTA of 4 warps per SM): /" ;;
    trio)   for l in 7 5; do
              if ASB_ADMM_LAYOUT=$l timeout 180 python -m pytest tests/test_gpu_ops.py -x -q -m gpu -k admm > gpurun_out/admm$l.log 2>&1; then
                ASB_ADMM_LAYOUT=$l timeout 300 python tools/kbench.py --only admm 2>&1 | grep admm_solve_x10 | sed "s/^/layout $l: /"
              else echo "layout $l tests FAILED"; tail -20 gpurun_out/admm$l.log; fi
            done ;;
    noalloc) timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/noalloc.log 2>&1; echo "noalloc (fresh process) rc=$?"; tail -4 gpurun_out/noalloc.log ;;
    ab)     timeout 900 python tools/admm_ab.py 14290 200 10 2 > gpurun_out/admm_ab.log 2>&1; echo "ab rc=$?"
            timeout 600 python tools/admm_ab.py 14290 200 1 2 >> gpurun_out/admm_ab.log 2>&1; cat gpurun_out/admm_ab.log ;;
    ncu_list) A="--grid 300 300 --frames 500 --comps 48 --splocs-iters 3 --steps 1 --warmup 1 --no-e2e --no-cpu-baseline"
            if timeout 600 python bench.py $A > gpurun_out/ncu_list_plain.json 2> gpurun_out/ncu_list_plain.err; then
              timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:asb:: -c 60000 --csv --log-file gpurun_out/ncu_launches.csv python bench.py $A > gpurun_out/ncu_list.log 2>&1; echo "ncu_list rc=$?"; tail -2 gpurun_out/ncu_list.log; wc -l gpurun_out/ncu_launches.csv
            else echo "plain run failed"; tail -5 gpurun_out/ncu_list_plain.err; fi ;;
    ncu_admm) if timeout 300 python tools/kbench.py --only admm > gpurun_out/kbench_admm_plain.log 2>&1; then
              grep admm_solve gpurun_out/kbench_admm_plain.log
              timeout 900 ncu --set full --clock-control none --import-source on -k regex:admm_tma_kernel -s 1 -c 1 -o gpurun_out/prof_admm_tma -f python tools/kbench.py --only admm > gpurun_out/ncu_admm.log 2>&1; echo "ncu_admm rc=$?"; tail -2 gpurun_out/ncu_admm.log; ls -la gpurun_out/*.ncu-rep
            else echo "plain kbench failed"; tail -5 gpurun_out/kbench_admm_plain.log; fi ;;
    bench)  timeout 1500 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; tail -c 4000 gpurun_out/bench_default.json; tail -5 gpurun_out/bench_default.err ;;
    benchN) N=${ASB_NGPU:-8}; timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "bench$N rc=$?"; tail -c 4000 gpurun_out/bench_${N}gpu.json; tail -5 gpurun_out/bench_${N}gpu.err ;;
    nccl2)  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tools/nccl_worker.py --mirror-classes > gpurun_out/nccl_worker_2gpu.log 2>&1; echo "nccl2 rc=$?"; tail -15 gpurun_out/nccl_worker_2gpu.log ;;
    ab5)    ASB_ADMM_LAYOUT=5 timeout 1200 python -m pytest tests/test_gpu_scale.py tests/test_gpu_ops.py tests/test_gpu_pipeline.py -x -q -m gpu > gpurun_out/layout5_tests.log 2>&1; echo "layout-5 scale+ops+pipeline rc=$?"; tail -3 gpurun_out/layout5_tests.log
            : > gpurun_out/admm_ab5.log
            timeout 600 python tools/admm_ab.py 14290 200 10 8 --variants 0 >> gpurun_out/admm_ab5.log 2>&1
            timeout 600 python tools/admm_ab.py 34834 200 10 4 --variants 0 >> gpurun_out/admm_ab5.log 2>&1
            timeout 600 python tools/admm_ab.py 200000 200 10 3 --variants 0 >> gpurun_out/admm_ab5.log 2>&1
            timeout 600 python tools/admm_ab.py 1000000 200 10 2 --variants 0 >> gpurun_out/admm_ab5.log 2>&1
            timeout 600 python tools/admm_ab.py 100000 136 7 3 --variants 0 >> gpurun_out/admm_ab5.log 2>&1
            timeout 600 python tools/admm_ab.py 50001 256 5 3 --variants 0 >> gpurun_out/admm_ab5.log 2>&1
            cat gpurun_out/admm_ab5.log ;;
    geoleaf) : > gpurun_out/geo_leaf.log; for l in 32 64 128 256; do timeout 600 python tools/kbench.py --geo 1000 1000 --leaf-size $l 2>&1 | grep geodesic_query >> gpurun_out/geo_leaf.log; done; cat gpurun_out/geo_leaf.log ;;
    ncu_geo) if timeout 600 python tools/kbench.py --geo 1000 1000 > gpurun_out/geo_plain.log 2>&1; then
              timeout 1500 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum --clock-control none --graph-profiling node --csv --log-file gpurun_out/ncu_geo.csv python tools/kbench.py --geo 1000 1000 > gpurun_out/ncu_geo.log 2>&1; echo "ncu_geo rc=$?"; tail -2 gpurun_out/ncu_geo.log; wc -l gpurun_out/ncu_geo.csv
            else echo "plain geo failed"; tail -5 gpurun_out/geo_plain.log; fi ;;
    geo1)   timeout 600 python tools/kbench.py --geo 1000 1000 2>&1 | grep geodesic_query | tee gpurun_out/geo1.log | cut -c1-600
            ASB_MF_SMALL=0 timeout 600 python tools/kbench.py --geo 1000 1000 2>&1 | grep geodesic_query | sed 's/^/ASB_MF_SMALL=0 /' | tee -a gpurun_out/geo1.log | cut -c1-600 ;;
    stream) timeout 600 python tools/kbench.py --only stream 2>&1 | tee gpurun_out/kbench_stream.log | tail -12 ;;
    traceN) N=${ASB_NGPU:-2}; for pr in 0 -1; do echo "== side priority $pr"; ASB_SIDE_PRIORITY=$pr ASB_TRACE_DEFLATE=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --splocs-iters 0 --steps 2 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/trace_${N}gpu.log 2>&1; echo "trace$N rc=$?"; grep -o "\[deflate trace.*\|\"phase_ms_per_step\": {[^}]*}" gpurun_out/trace_${N}gpu.log | tail -3; done ;;
    deflN)  N=${ASB_NGPU:-8}; for cfg in "ASB_DEFLATE_PREDICT=1 ASB_MAIN_PRIORITY=0" "ASB_DEFLATE_PREDICT=1 ASB_MAIN_PRIORITY=-1" "ASB_DEFLATE_PREDICT=0 ASB_MAIN_PRIORITY=0"; do
              env $cfg timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $N --splocs-iters 0 --steps 3 --warmup 2 --no-e2e --no-cpu-baseline > gpurun_out/defl_${N}gpu.log 2>&1
              echo "$cfg rc=$? $(grep -o '"phase_ms_per_step": {[^}]*}' gpurun_out/defl_${N}gpu.log | tail -1) $(grep -o "'predicted_picks': [0-9]*" gpurun_out/defl_${N}gpu.log | tail -1)"; done ;;
    strace) for pr in 0 -1; do echo "== side priority $pr"; ASB_SIDE_PRIORITY=$pr ASB_TRACE_SPLOCS=1 timeout 600 python bench.py --splocs-iters 12 --steps 1 --warmup 1 --no-e2e --no-cpu-baseline 2>&1 | grep -o "\[splocs trace.*\|\"phase_ms_per_step\": {[^}]*}" | tail -2; done ;;
    cache)  for cr in 12; do ASB_GEO_CACHE_ROWS=$cr timeout 900 python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/bench_cache$cr.json 2> gpurun_out/bench_cache$cr.err; echo "cache_rows=$cr rc=$? $(grep -o '"value": [0-9.]*' gpurun_out/bench_cache$cr.json | head -1) $(grep -o '"phase_ms_per_step": {[^}]*}' gpurun_out/bench_cache$cr.json) $(grep -o '"geodesic": {[^}]*}' gpurun_out/bench_cache$cr.json)"; done ;;
    final)  timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/tests_final.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/tests_final.log
            timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
            ( time timeout 1500 python bench.py ) > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench default rc=$?"; tail -c 1500 gpurun_out/bench_default.json; tail -6 gpurun_out/bench_default.err
            ( time timeout 900 python bench.py --impl reference --steps 1 --warmup 0 ) > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "bench reference rc=$?"; tail -c 1200 gpurun_out/bench_reference.json; tail -4 gpurun_out/bench_reference.err ;;
    f1000)  timeout 900 python -m pytest tests/test_gpu_scale.py -q -m gpu -k f1000 -rs > gpurun_out/scale_f1000.log 2>&1; echo "scale f1000 rc=$?"; tail -3 gpurun_out/scale_f1000.log
            c=armadillo_splocs_f1000; timeout 600 python bench.py --config $c --steps 3 --warmup 3 > gpurun_out/bench_$c.json 2> gpurun_out/bench_$c.err; echo "$c rc=$?"; tail -c 1500 gpurun_out/bench_$c.json; tail -3 gpurun_out/bench_$c.err ;;
    *) echo "unknown step $step" ;;
  esac
done
