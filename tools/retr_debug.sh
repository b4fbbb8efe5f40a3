#!/bin/bash
for cfg in "REIDB200_KNN_SELECT=0" "REIDB200_KNN_SELECT=1" "REIDB200_KNN_SELECT=1 REIDB200_SEL_DEBUG=16" "REIDB200_KNN_SELECT=1 REIDB200_SEL_DEBUG=31" "REIDB200_KNN_SELECT=1 REIDB200_SEL_DEBUG=4" "REIDB200_KNN_SELECT=1 REIDB200_SEL_DEBUG=2"; do
  echo -n "$cfg  "
  env $cfg timeout 300 python bench.py --steps 3 --warmup 3 --no-assoc --no-cpu --no-sharded-rerank --retrieval-gallery 1000000 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['retrieval']
print(round(r['ms_per_search'],1), r['kernel_ms'])"
done
