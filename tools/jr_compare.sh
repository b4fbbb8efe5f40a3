#!/bin/bash
# the evaluation row kernel (jaccard_eval_rows_kernel) against the general one it replaces on the fused path
# (REIDB200_JR_OLD=1): parity tests first, then the headline step under both and under the timing switches
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_rerank.py tests/test_gpu_hard_eval.py tests/test_gpu_fullsize.py tests/test_gpu_shard.py tests/test_gpu_select.py -q -m gpu --timeout=600 2>&1 | tail -8
for v in "REIDB200_JR_OLD=1" "REIDB200_JE_DEBUG=0" $EXTRA; do
  echo "== $v"
  env $v timeout 300 python bench.py --no-assoc --no-cpu --no-retrieval --no-cpu-rows --no-sharded-rerank --steps 20 --warmup 3 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
sh = d['kernel_share_of_step']; ev = d['ms_per_step_with_kernel_events']
print('ms_per_step', round(d['ms_per_step'], 4), 'easy', round(d['survey_recipe']['ms_per_step'], 4), 'jaccard_rows_ms', round(sh['jaccard_rows_kernel'] * ev, 4), 'gemm_ms', round(sh['gemm_split_tc_kernel'] * ev, 4))
print('result', d['result'], d['survey_recipe']['result'])
print({k: round(v * ev * 1000, 1) for k, v in sh.items()})
"
done 2>&1 | tee gpurun_out/jr_compare.log
