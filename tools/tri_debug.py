"""Diagnostic of triangle mode (REIDB200_TRI): runs the fused re-ranked evaluation in both modes on POISONED workspaces
(every scratch byte 0xFF before the call, so that a read of a place the call itself did not write shows) and reports
where the outputs differ."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import reidb200
from reidb200 import ops, synth, _lib as L

orig_ws = L.workspace
poison = [False]
def ws(nbytes, device):
    t = orig_ws(nbytes, device)
    if poison[0]:
        t.fill_(0xFF)
    return t
L.workspace = ws
ops.L.workspace = ws

def run(N, D, k1, k2, seed):
    Q = N // 6; G = N - Q
    d = synth.market_like(Q=Q, G=G, D=D, n_ids=max(N // 25, 8), seed=seed, noise="hard")
    merged = torch.from_numpy(np.concatenate([d["gf"], d["qf"]], 0)).cuda()
    lab = [torch.from_numpy(d[k].astype(np.int32)).cuda() for k in ("ql", "qc", "gl", "gc")]
    outs = []
    for p in (False, True):
        poison[0] = p
        for mode in ("0", "2", "0", "2"):
            os.environ["REIDB200_TRI"] = mode
            cmc, ap, dist, ap_q, first = ops.rerank_eval(merged, G, *lab, k1, k2, 0.3, want_dist=True, per_query=True)
            torch.cuda.synchronize()
            outs.append((p, mode, cmc.clone(), ap.clone(), dist.clone(), ap_q.clone(), first.clone()))
    os.environ.pop("REIDB200_TRI")
    poison[0] = False
    J, rank = ops.jaccard(merged, k1, k2, want_rank=True)
    d2 = ops.jaccard(merged, k1, k2, False, 0.3, rows=(G, Q), cols=(0, G))
    ref = outs[0]
    print("== N=%d D=%d k1=%d k2=%d G=%d Q=%d" % (N, D, k1, k2, G, Q))
    print("two-call vs first:", int((d2 != ref[4]).sum()), "nan in two-call", int(torch.isnan(d2).sum()))
    for o in outs:
        diff = (o[4] != ref[4]) & ~(torch.isnan(o[4]) & torch.isnan(ref[4]))
        n = int(diff.sum())
        msg = "poison=%d mode=%s: dist differs in %d places, nan %d, cmc equal %s, ap %.12f, ap_q differs %d, first differs %d" % (
            o[0], o[1], n, int(torch.isnan(o[4]).sum()), bool(torch.equal(o[2], ref[2])), float(o[3][0]), int((o[5] != ref[5]).sum()), int((o[6] != ref[6]).sum()))
        print(msg)
        if n:
            idx = diff.nonzero()[:12].cpu().numpy()
            rows = np.unique(diff.nonzero()[:, 0].cpu().numpy())
            cols = np.unique(diff.nonzero()[:, 1].cpu().numpy())
            print("   rows (query index):", rows[:20], "n rows", len(rows), " cols:", cols[:20], "n cols", len(cols))
            for r, c in idx[:6]:
                print("   (%d,%d): %.9g vs %.9g" % (r, c, float(o[4][r, c]), float(ref[4][r, c])))

cfgs = [(2305, 48, 16, 1, 2306), (2311, 32, 12, 3, 2312), (3000, 128, 20, 6, 3001), (2305, 48, 16, 6, 2306), (2305, 64, 16, 1, 2306), (2304, 48, 16, 1, 2306)]
for c in cfgs:
    try:
        run(*c)
    except Exception as e:
        print("ERROR", c, repr(e)[:300])
