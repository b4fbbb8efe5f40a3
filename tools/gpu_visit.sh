#!/bin/bash
# One GPU visit: every GPU test (full report, not -x), smoke, then the bench lines.  Usage: tools/gpu_visit.sh <tag> [bench args]
tag=${1:-v}; shift
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu_$tag.txt 2>&1
timeout 1500 python -m pytest tests -q -m gpu --timeout=900 > gpurun_out/t_$tag.log 2>&1; echo "tests rc=$?" | tee gpurun_out/summary_$tag.txt
grep -E 'passed|failed|FAILED|ERROR|full-size' gpurun_out/t_$tag.log | tail -30
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke_$tag.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/summary_$tag.txt
tail -3 gpurun_out/smoke_$tag.log
if [ "$NO_BENCH" != "1" ]; then
  ( time timeout 420 python bench.py "$@" > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err ) 2>&1 | tail -3; echo "bench rc=$?" | tee -a gpurun_out/summary_$tag.txt
  tail -5 gpurun_out/bench_$tag.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_$tag.json").read().strip().splitlines()[-1])
    keep={k:d[k] for k in ("value","ms_per_step","e2e","roofline","kernel_share_of_step","survey_recipe","result","cpu_baseline","retrieval_ms_per_search") if k in d}
    print(json.dumps(keep)[:3000])
    if "assoc" in d: print(json.dumps(d["assoc"])[:2500])
    if "cpu_rows" in d: print(json.dumps(d["cpu_rows"])[:2500])
except Exception as e:
    print("no bench line", e)
PY
fi
if [ "$WITH_REF" = "1" ]; then
  ( time timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_$tag.json 2> gpurun_out/bench_ref_$tag.err ) 2>&1 | tail -3
  cat gpurun_out/bench_ref_$tag.json | cut -c1-1500
fi
