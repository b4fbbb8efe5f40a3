#!/bin/bash
mkdir -p gpurun_out
timeout 700 python -m pytest tests -q -m gpu --timeout=300 -x > gpurun_out/t25.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/t25.log
timeout 900 python bench.py --steps 10 --warmup 3 --no-assoc --no-cpu --no-retrieval > gpurun_out/bench25.json 2> gpurun_out/bench25.err; echo "bench rc=$?"
tail -5 gpurun_out/bench25.err; python -c "
import json;d=json.load(open('gpurun_out/bench25.json'));print(d['value'],d['ms_per_step'],d['e2e'])"
