#!/bin/bash
# crop kernel variants side by side (bench.bench_assoc under the switches): v4 (scalar fp32) vs v4p (packed f32x2, default)
for v in "REIDB200_CROP_V4=1" "REIDB200_CROP_DEFAULT=1"; do
  echo "== $v"
  env $v timeout 600 python - <<'PY'
import json, sys
sys.path.insert(0, ".")
import torch, bench
r = bench.bench_assoc(5, 3)
for k, v in r.items():
    print(k, json.dumps({x: round(v[x], 2) for x in ("us_per_frame", "us_per_frame_cuda_graph", "e2e_us_per_frame", "crop_kernel_us", "assoc_kernel_us", "crop_gbs")}))
PY
done
