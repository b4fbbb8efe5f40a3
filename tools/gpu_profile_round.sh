#!/bin/bash
# Evidence visit: GPU tests + smoke, full bench (both arms), ncu launch list of the bench command, ncu --set full of one step.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu --timeout=600 > gpurun_out/tests_gpu.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/tests_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 1200 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_full.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-retrieval --no-assoc > gpurun_out/ncu_launch.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -s 12 -c 12 -f -o gpurun_out/r01_step_full_c python tools/step_once_fused.py 2 > gpurun_out/ncu_step.log 2>&1; echo "ncu step rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:dbscan -s 4 -c 4 -f -o gpurun_out/r01_dbscan_full python tools/dbscan_once.py 2 > gpurun_out/ncu_dbscan.log 2>&1; echo "ncu dbscan rc=$?"
