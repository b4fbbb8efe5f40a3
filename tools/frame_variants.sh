#!/bin/bash
# crop / association kernel variants side by side (bench.bench_assoc under the opt-in switches)
for v in "REIDB200_CROP_V3=1 REIDB200_ASSOC_V1=1" "REIDB200_NEW=1"; do
  echo "== $v"
  env $v timeout 600 python - <<'PY'
import json, sys
sys.path.insert(0, ".")
import torch, bench
r = bench.bench_assoc(5, 3)
for k, v in r.items():
    print(k, json.dumps({x: v[x] for x in v if x != "match_e2e_us"}))
PY
done
