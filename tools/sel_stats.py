"""Diagnostics: candidate-list statistics of the selection fused into the GEMM epilogue (REIDB200_SEL_STATS=1) and the
time of one re-ranked evaluation step, at the full Market shape."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["REIDB200_SEL_STATS"] = "1"
import numpy as np, torch
import reidb200
from reidb200 import ops, synth, _lib as L
hard = len(sys.argv) > 1 and sys.argv[1] == "hard"
d = synth.market_like(seed=0, noise=2.5 if hard else 0.6)
G, Q = len(d["gl"]), len(d["ql"])
merged = torch.from_numpy(np.concatenate([d["gf"], d["qf"]], 0)).cuda()
lab = [torch.from_numpy(d[k].astype(np.int32)).cuda() for k in ("ql", "qc", "gl", "gc")]
for it in range(2):
    t0 = time.perf_counter()
    cmc, ap = ops.rerank_eval(merged, G, *lab, 20, 6, 0.3)
    torch.cuda.synchronize()
    print("step %d: %.2f ms  rank1 %.6f mAP %.6f" % (it, (time.perf_counter() - t0) * 1e3, float(cmc[0]), float(ap[0])))
del os.environ["REIDB200_SEL_STATS"]
L.profile_enable(True)
for _ in range(5):
    ops.rerank_eval(merged, G, *lab, 20, 6, 0.3)
torch.cuda.synchronize()
p = L.profile_read()
print({k: round(v[1] / v[0], 4) for k, v in sorted(p.items(), key=lambda kv: -kv[1][1])})
D, I = ops.knn(merged[G:G + 1000], merged[:G], 100, "sql2")
torch.cuda.synchronize()
