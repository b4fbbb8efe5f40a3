#!/bin/bash
mkdir -p gpurun_out
bash tools/gpu_check.sh 2>&1 | tail -40
timeout 900 python bench.py --steps 10 --warmup 3 --no-retrieval --no-cpu > gpurun_out/bench5.json 2> gpurun_out/bench5.err; echo "bench rc=$?"
tail -5 gpurun_out/bench5.err; python -c "
import json;d=json.load(open('gpurun_out/bench5.json'));print(d['value'],d['ms_per_step'],d['roofline']['frac'],d['kernel_share_of_step']);print(json.dumps(d['assoc']))"
