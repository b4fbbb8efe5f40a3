"""Two passes of the Market-shaped re-ranked evaluation step (for ncu captures of single kernels)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import reidb200
from reidb200 import ops, synth
import bench

d = synth.market_like(Q=bench.Q_FULL, G=bench.G_FULL, D=bench.D_FULL, n_ids=bench.IDS_FULL, seed=0)
dev = torch.device("cuda", 0)
merged = torch.from_numpy(np.concatenate([d["gf"], d["qf"]], 0)).to(dev)
lab = {k: torch.from_numpy(d[k].astype(np.int32)).to(dev) for k in ("ql", "qc", "gl", "gc")}
G, Q = bench.G_FULL, bench.Q_FULL
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    dist = ops.jaccard(merged, 20, 6, False, 0.3, rows=(G, Q), cols=(0, G))
    cmc, ap = ops.eval_cmc_map(dist, lab["ql"], lab["qc"], lab["gl"], lab["gc"], descending=False)
torch.cuda.synchronize()
print("ok", float(ap[0]))
