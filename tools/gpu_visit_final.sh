#!/bin/bash
# Re-entry validation: smoke, every GPU test, default bench, reference arm.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke29.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke29.log
timeout 1200 python -m pytest tests -q -m gpu --timeout=300 > gpurun_out/t29.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/t29.log
timeout 900 python bench.py > gpurun_out/bench29.json 2> gpurun_out/bench29.err; echo "bench rc=$?"
tail -3 gpurun_out/bench29.err; python -c "
import json;d=json.load(open('gpurun_out/bench29.json'));print(d['value'],d['ms_per_step'],d['e2e'],d['roofline']['frac'])"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref29.json 2> gpurun_out/bench_ref29.err; echo "ref rc=$?"; cat gpurun_out/bench_ref29.json
# launch list of the bench command (after it has exited 0 without ncu above)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches29.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-retrieval --no-assoc > gpurun_out/ncu_launch29.log 2>&1; echo "ncu launches rc=$?"
