"""End-to-end pipelining depth experiment: evaluate.RerankedEvalStream with depth 2 / 3 / 4 (sets in flight = depth), the
same pinned host inputs as bench.py's e2e leg.  Prints ms per evaluation and pairs/s for each depth, plus the H2D time of
one set alone."""
import os, sys, time, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import reidb200
from reidb200 import evaluate
import bench

d = bench.market_set(0, hard=True)
G, Q = bench.G_FULL, d["qf"].shape[0]
merged_host = torch.from_numpy(np.concatenate([d["gf"], d["qf"]], 0)).pin_memory()
lab_host = {k: torch.from_numpy(d[k].astype(np.int32)).pin_memory() for k in ("ql", "qc", "gl", "gc")}
qf_h, gf_h = merged_host[G:], merged_host[:G]
dev = torch.device("cuda", 0)
buf = torch.empty_like(merged_host, device=dev)
for _ in range(3):
    buf.copy_(merged_host, non_blocking=True)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20):
    buf.copy_(merged_host, non_blocking=True)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 20
print("H2D of one set alone: %.3f ms = %.1f GB/s" % (dt * 1e3, merged_host.numel() * 4 / dt / 1e9))
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
for depth in (2, 3, 4, 2, 3):
    ev = evaluate.RerankedEvalStream(depth=depth)
    def run(n):
        pend = collections.deque()
        out = None
        for _ in range(n):
            pend.append(ev.submit(qf_h, lab_host["ql"], lab_host["qc"], gf_h, lab_host["gl"], lab_host["gc"], bench.K1, bench.K2, bench.LAMBDA))
            if len(pend) >= depth:
                out = ev.result(pend.popleft())
        while pend:
            out = ev.result(pend.popleft())
        return out
    run(depth + 2)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    cmc, ap = run(steps)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / steps
    print("depth %d: %.4f ms per evaluation = %.2f G pairs/s  (mAP %.6f)" % (depth, dt * 1e3, Q * G / dt / 1e9, ap))
    del ev
    evaluate.release_plans()
    torch.cuda.empty_cache()
