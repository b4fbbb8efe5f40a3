"""Two passes of the fused re-ranked evaluation step (ops.rerank_eval) for ncu captures of single kernels."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import reidb200
from reidb200 import ops, synth
import bench

# argv[2] = 'easy' for SURVEY 8d's recipe (noise 0.6); default = the hard data of the headline
d = bench.market_set(0, hard=not (len(sys.argv) > 2 and sys.argv[2] == 'easy'))
dev = torch.device("cuda", 0)
merged = torch.from_numpy(np.concatenate([d["gf"], d["qf"]], 0)).to(dev)
lab = {k: torch.from_numpy(d[k].astype(np.int32)).to(dev) for k in ("ql", "qc", "gl", "gc")}
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    cmc, ap = ops.rerank_eval(merged, bench.G_FULL, lab["ql"], lab["qc"], lab["gl"], lab["gc"], 20, 6, 0.3)
torch.cuda.synchronize()
print("ok", float(ap[0]))
