#!/bin/bash
# timing-only experiments on the selection epilogue (REIDB200_SEL_DEBUG bits; results are wrong when set)
for dbg in 0 16 1 2 4 8 31; do
  echo -n "SEL_DEBUG=$dbg  "
  REIDB200_SEL_DEBUG=$dbg timeout 120 python - <<'PY' 2>&1 | tail -1
import os, sys
sys.path.insert(0, ".")
import numpy as np, torch
import reidb200
from reidb200 import ops, synth, _lib as L
d = synth.market_like(seed=0, noise=2.5)
G = len(d["gl"])
merged = torch.from_numpy(np.concatenate([d["gf"], d["qf"]], 0)).cuda()
lab = [torch.from_numpy(d[k].astype(np.int32)).cuda() for k in ("ql", "qc", "gl", "gc")]
try:
    for _ in range(2): ops.rerank_eval(merged, G, *lab, 20, 6, 0.3)
    L.profile_enable(True)
    for _ in range(5): ops.rerank_eval(merged, G, *lab, 20, 6, 0.3)
    torch.cuda.synchronize()
    p = L.profile_read()
    print({k: round(v[1] / v[0], 3) for k, v in p.items() if k in ("gemm_split_tc_kernel", "select_lists_kernel", "sel_redo_kernel")})
except Exception as e:
    print("failed:", str(e)[:100])
PY
done
