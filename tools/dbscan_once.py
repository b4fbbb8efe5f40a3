"""The Market-shaped Jaccard matrix + DBSCAN passes on it (for ncu captures and quick timings of the N4 kernels)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import reidb200
from reidb200 import _lib as L, ops, synth
import bench

d = synth.market_like(Q=bench.Q_FULL, G=bench.G_FULL, D=bench.D_FULL, n_ids=bench.IDS_FULL, seed=0)
dev = torch.device("cuda", 0)
merged = torch.from_numpy(np.concatenate([d["gf"], d["qf"]], 0)).to(dev)
J = ops.jaccard(merged, 20, 6)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
L.profile_enable(True)
for _ in range(n):
    labels, ncl, core = ops.dbscan_precomputed(J, 0.5, 7)
torch.cuda.synchronize()
print({k: round(v[1] / v[0], 4) for k, v in L.profile_read().items()})
L.profile_enable(False)
hits = (J <= 0.5).sum(1)
print("ok clusters", int(ncl.item()), "core", int(core.sum()), "hits/row mean", float(hits.float().mean()), "max", int(hits.max()))
