#!/bin/bash
# same-box A/B of the headline step under environment switches: bash tools/ab_env.sh "VAR=1" "OTHER=1" ... (first arm: none)
mkdir -p gpurun_out
for rep in 1 2; do
for v in "REIDB200_NONE=1" "$@"; do
  echo "== $v"
  env $v timeout 300 python bench.py --no-assoc --no-cpu --no-retrieval --no-cpu-rows --no-sharded-rerank --steps 30 --warmup 3 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('ms_per_step', round(d['ms_per_step'], 4), 'easy', round(d['survey_recipe']['ms_per_step'], 4), 'stream launches', round(d['ms_per_step_stream_launches'], 4), 'e2e', round(d['e2e']['value'] / 1e9, 2))
"
done
done 2>&1 | tee gpurun_out/ab_env.log
