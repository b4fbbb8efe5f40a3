#!/bin/bash
# retrieval (configs[4]: 10 k x 4 M x 2048, top-100) with larger distance blocks per GEMM + select pass
# (REIDB200_KNN_BLOCK_FLOATS; the default is 2^29 floats = 2 GiB)
for v in 536870912 2147483648 4294967296 8589934592; do
  echo "== REIDB200_KNN_BLOCK_FLOATS=$v"
  REIDB200_KNN_BLOCK_FLOATS=$v timeout 400 python bench.py --steps 3 --warmup 3 --no-assoc --no-cpu --no-cpu-rows --no-sharded-rerank 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['retrieval']
print(json.dumps({k:r[k] for k in ('ms_per_search','pairs_per_s','gemm_frac_of_peak_div3','kernel_ms','launches','checksum')}))"
done 2>&1 | tee gpurun_out/retr_blocks.log
