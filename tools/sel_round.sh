#!/bin/bash
timeout 300 python -m pytest tests/test_gpu_select.py tests/test_gpu_dist.py -q -m gpu 2>&1 | tail -3
for dbg in 0 16; do echo -n "market SEL_DEBUG=$dbg  "; REIDB200_RERANK_SELECT=1 REIDB200_SEL_DEBUG=$dbg timeout 120 python - <<'PY' 2>&1 | tail -1
import os, sys
sys.path.insert(0, ".")
import numpy as np, torch
import reidb200
from reidb200 import ops, synth, _lib as L
d = synth.market_like(seed=0, noise=2.5)
G = len(d["gl"])
merged = torch.from_numpy(np.concatenate([d["gf"], d["qf"]], 0)).cuda()
lab = [torch.from_numpy(d[k].astype(np.int32)).cuda() for k in ("ql", "qc", "gl", "gc")]
for _ in range(2): ops.rerank_eval(merged, G, *lab, 20, 6, 0.3)
L.profile_enable(True)
for _ in range(5): ops.rerank_eval(merged, G, *lab, 20, 6, 0.3)
torch.cuda.synchronize()
p = L.profile_read()
print({k: round(v[1] / v[0], 3) for k, v in p.items() if k in ("gemm_split_tc_kernel", "select_lists_kernel", "sel_redo_kernel", "expand_v_kernel")})
PY
done
for cfg in "REIDB200_KNN_SELECT=1" "REIDB200_KNN_SELECT=1 REIDB200_SEL_DEBUG=16"; do
  echo -n "retr1M $cfg  "
  env $cfg timeout 300 python bench.py --steps 3 --warmup 3 --no-assoc --no-cpu --no-sharded-rerank --retrieval-gallery 1000000 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['retrieval']
print(round(r['ms_per_search'],1), r['kernel_ms'])"
done
