#!/bin/bash
# Re-entry validation: smoke, every GPU test, default bench, reference arm.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke26.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke26.log
timeout 1200 python -m pytest tests -q -m gpu --timeout=300 > gpurun_out/t26.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/t26.log
timeout 900 python bench.py > gpurun_out/bench26.json 2> gpurun_out/bench26.err; echo "bench rc=$?"
tail -3 gpurun_out/bench26.err; python -c "
import json;d=json.load(open('gpurun_out/bench26.json'));print(d['value'],d['ms_per_step'],d['e2e'],d['roofline']['frac'])"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref26.json 2> gpurun_out/bench_ref26.err; echo "ref rc=$?"; cat gpurun_out/bench_ref26.json
