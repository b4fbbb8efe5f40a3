#!/bin/bash
# frame context variants (REIDB200_FRAME_MODE: bit 0 = cost blocks written straight to mapped pinned memory, bit 1 = the
# per-frame tables go up on a side stream) through bench.bench_assoc; and the frame-context parity test in each mode
mkdir -p gpurun_out
for m in 0 1 2 3; do
  echo "== REIDB200_FRAME_MODE=$m"
  REIDB200_FRAME_MODE=$m timeout 300 python -m pytest tests/test_gpu_assoc.py -q -m gpu -k "frame_context" 2>&1 | tail -1
  REIDB200_FRAME_MODE=$m timeout 600 python - <<'PY'
import json, sys
sys.path.insert(0, ".")
import torch, bench
for rep in range(2):
    r = bench.bench_assoc(5, 3)
    for k, v in r.items():
        print(k, json.dumps({x: round(v[x], 2) for x in ("us_per_frame", "us_per_frame_cuda_graph", "e2e_us_per_frame", "e2e_us_per_frame_python_marshalling", "crop_kernel_us", "assoc_kernel_us")}))
PY
done 2>&1 | tee gpurun_out/frame_modes.log
