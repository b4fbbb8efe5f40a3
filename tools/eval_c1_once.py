"""Per-kernel time of one plain Market-shaped evaluate_all (configs[0]) through shard.evaluate_all_sharded."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import reidb200  # noqa: F401
from reidb200 import _lib as L, shard, synth

dev = torch.device("cuda", 0)
d = synth.market_like(seed=0)
a = [torch.from_numpy(d[k]).to(dev) for k in ("qf", "ql", "qc", "gf", "gl", "gc")]
a[1], a[2], a[4], a[5] = (x.to(torch.int32) for x in (a[1], a[2], a[4], a[5]))
for _ in range(3):
    shard.evaluate_all_sharded(*a, device=dev)
torch.cuda.synchronize()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(n):
    cmc, ap = shard.evaluate_all_sharded(*a, device=dev)
e1.record()
torch.cuda.synchronize()
print("ms per evaluate_all: %.4f" % (e0.elapsed_time(e1) / n))
L.profile_enable(True)
for _ in range(n):
    shard.evaluate_all_sharded(*a, device=dev)
torch.cuda.synchronize()
for k, (c, ms) in sorted(L.profile_read().items(), key=lambda kv: -kv[1][1]):
    print("%-32s %3d launches  %.4f ms each" % (k, c // n, ms / c))
L.profile_enable(False)
print("rank1 %.6f mAP %.6f" % (float(cmc[0]), float(ap)))
