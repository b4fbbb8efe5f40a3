#!/bin/bash
mkdir -p gpurun_out
for t in tests/test_gpu_crop.py tests/test_gpu_assoc.py; do
  n=$(basename $t .py)
  timeout 600 python -m pytest $t -q -m gpu --timeout=300 -x > gpurun_out/$n.log 2>&1; echo "$n rc=$?" | tee -a gpurun_out/summary3.txt
  tail -25 gpurun_out/$n.log
done
timeout 900 python bench.py --steps 10 --warmup 3 --no-retrieval --no-cpu > gpurun_out/bench3.json 2> gpurun_out/bench3.err; echo "bench rc=$?" | tee -a gpurun_out/summary3.txt
tail -5 gpurun_out/bench3.err; python -c "
import json;d=json.load(open('gpurun_out/bench3.json'));print(json.dumps(d['assoc'],indent=1))"
