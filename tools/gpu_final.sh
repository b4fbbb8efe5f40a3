#!/bin/bash
# lean end-of-round visit: every GPU test (-x like the driver), smoke, the default bench line
tag=${1:-final}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu --timeout=600 > gpurun_out/t_$tag.log 2>&1; echo "tests rc=$?" | tee gpurun_out/summary_$tag.txt
tail -2 gpurun_out/t_$tag.log | cut -c1-300
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke_$tag.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/summary_$tag.txt
tail -2 gpurun_out/smoke_$tag.log | cut -c1-300
( time timeout 420 python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err ) 2>&1 | tail -3; echo "bench done" | tee -a gpurun_out/summary_$tag.txt
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_$tag.json").read().strip().splitlines()[-1])
keep={k:d[k] for k in ("value","ms_per_step","e2e","roofline","kernel_share_of_step","survey_recipe","result","retrieval_ms_per_search","gpu_launches","clocks") if k in d}
print(json.dumps(keep)[:2600])
print(json.dumps(d["assoc"]["mot20"])[:600])
PY
