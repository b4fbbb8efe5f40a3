"""Where one row of jaccard_eval_rows_kernel spends its time: cycles per phase summed by thread 0 of every CTA
(REIDB200_JE_DEBUG=16), on the hard Market-shaped set.  usage: python tools/je_phases.py [easy]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["REIDB200_JE_DEBUG"] = str(16 | int(os.environ.get("JE_EXTRA", "0")))
import numpy as np, torch
import reidb200
from reidb200 import ops
import bench

d = bench.market_set(0, hard=not (len(sys.argv) > 1 and sys.argv[1] == "easy"))
dev = torch.device("cuda", 0)
G, Q = bench.G_FULL, bench.Q_FULL
merged = torch.from_numpy(np.concatenate([d["gf"], d["qf"]], 0)).to(dev)
lab = [torch.from_numpy(d[k].astype(np.int32)).to(dev) for k in ("ql", "qc", "gl", "gc")]
rd, ri = ops.knn(merged, merged, 20, "sql2", idx64=False)
ld = (G + 31) // 32 * 32
blend = torch.empty((Q, ld), dtype=torch.float32, device=dev)
ops.distmat(merged[G:], merged[:G], "sql2", out=blend[:, :G])
for rep in range(2):
    _, hist, ap_q, first = ops.rerank_rows_from_rank(merged, ri, rd, 20, 6, (G, Q), (0, G), 0.3, blend, want_out=False, labels=lab)
torch.cuda.synchronize()
st = ops._last_status.cpu().numpy()[8:15].astype(np.float64) * 16 / Q
names = ["loads+entries+zero", "scan | label lists", "scatter", "goods+sort", "column test", "candidates", "AP tail"]
print("mean cycles per row (thread 0's view), mAP %.6f" % float(ap_q.mean()))
for n, v in zip(names, st):
    print("  %-22s %8.0f cycles  %5.1f %%" % (n, v, 100 * v / st.sum()))
print("  total %.0f cycles = %.2f us at 1.9 GHz" % (st.sum(), st.sum() / 1900))
raw = ops._last_status.cpu().numpy()
print("  candidates per row: mean %.0f max %d, rows over the staging capacity %d; scatter terms per row: mean %.0f max %d; V_qe non-zeros per row %.0f"
      % (raw[16] / Q, raw[17], raw[18], raw[19] * 16 / Q, raw[20], raw[21] / Q))
