#!/bin/bash
# retrieval (configs[4]) with the k-NN selected in the GEMM epilogue vs the distance-block path
for v in 0 1; do
  echo "== REIDB200_KNN_SELECT=$v"
  REIDB200_KNN_SELECT=$v timeout 400 python bench.py --steps 3 --warmup 3 --no-assoc --no-cpu --no-sharded-rerank 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['retrieval']
print(json.dumps({k:r[k] for k in ('ms_per_search','pairs_per_s','e2e_pairs_per_s','gemm_frac_of_peak_div3','kernel_ms','checksum')}))
print('step ms', d['ms_per_step'], d['roofline']['frac'])"
done
