#!/bin/bash
# same-box A/B of the headline step with per-kernel microseconds: bash tools/ab_kernels.sh "VAR=1" ...
mkdir -p gpurun_out
for v in "REIDB200_NONE=1" "$@"; do
  echo "== $v"
  env $v timeout 300 python bench.py --no-assoc --no-cpu --no-retrieval --no-cpu-rows --no-sharded-rerank --steps 20 --warmup 3 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
sh = d['kernel_share_of_step']; ev = d['ms_per_step_with_kernel_events']
print('ms_per_step', round(d['ms_per_step'], 4), 'easy', round(d['survey_recipe']['ms_per_step'], 4))
print({k: round(v * ev * 1000, 1) for k, v in sh.items()})
"
done 2>&1 | tee gpurun_out/ab_kernels.log
