#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_lap.py -q -m gpu --timeout=120 -x > gpurun_out/t24.log 2>&1; echo "tests rc=$?"; tail -25 gpurun_out/t24.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-retrieval > gpurun_out/bench24.json 2> gpurun_out/bench24.err; echo "bench rc=$?"; tail -5 gpurun_out/bench24.err
python -c "
import json;d=json.load(open('gpurun_out/bench24.json'));print(d['ms_per_step']);[print(k,v['e2e_us_per_frame'],v['match_e2e_us']) for k,v in d['assoc'].items()]"
