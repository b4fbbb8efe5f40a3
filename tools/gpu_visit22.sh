#!/bin/bash
mkdir -p gpurun_out
timeout 700 python -m pytest tests -q -m gpu --timeout=300 -x > gpurun_out/t22.log 2>&1; echo "tests rc=$?"; tail -30 gpurun_out/t22.log
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu --no-retrieval > gpurun_out/bench22.json 2> gpurun_out/bench22.err; echo "bench rc=$?"
tail -5 gpurun_out/bench22.err; python -c "
import json;d=json.load(open('gpurun_out/bench22.json'));print(d['value'],d['ms_per_step'],d['ms_per_step_stream_launches'],d['e2e']['value']);print(json.dumps(d['next_rows']['n4'],indent=1))"
