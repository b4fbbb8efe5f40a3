#!/bin/bash
# top-k load batching (REIDB200_TOPK_PIPE=0 = one load round per 32 chunks): the tests of every caller, then same-box A/B of
# the headline step and of the 10 k x 4 M retrieval
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_zy_triangle.py tests/test_gpu_select.py tests/test_gpu_rerank.py tests/test_gpu_dist.py tests/test_gpu_shard.py tests/test_gpu_hard_eval.py -q -m gpu --timeout=300 > gpurun_out/t_pipe.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/t_pipe.log | cut -c1-300
timeout 200 bash tools/ab_kernels.sh REIDB200_TOPK_PIPE=0 2>&1 | tail -6
cp gpurun_out/ab_kernels.log gpurun_out/ab_kernels_pipe.log
for v in REIDB200_NONE=1 REIDB200_TOPK_PIPE=0; do
  echo "== retrieval $v"
  env $v timeout 200 python bench.py --no-assoc --no-cpu --no-cpu-rows --no-sharded-rerank --steps 3 --warmup 3 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); r = d['retrieval']
print('ms_per_search', r['ms_per_search'], r['kernel_ms'], 'checksum', r['checksum'])
"
done 2>&1 | tee gpurun_out/retr_pipe.log
