"""Diagnostic: run the fused re-ranked evaluation in both modes and compare the WORKSPACE contents (matrix, chunk words,
initial ranking, V rows) -- offsets replicate carve_jaccard / carve_sparse in csrc/rerank.cu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import reidb200
from reidb200 import ops, synth, _lib as L

last = [None]
orig_ws = L.workspace
def ws(nbytes, device):
    t = orig_ws(nbytes, device); t.fill_(0xFF); last[0] = t; return t
L.workspace = ws
ops.L.workspace = ws

def al(x, a=256): return (x + a - 1) // a * a

def layout(N, D, k1, k2):
    dp = (D + 63) // 64 * 64
    half = int(np.rint(k1 / 2.0)); wh = min(half + 1, k1); cap1 = k1 + k1 * wh
    lds = al(N, 32); ldcr = (lds + 31) // 32
    off = 0; o = {}
    def take(name, nbytes):
        nonlocal off
        off = al(off); o[name] = off; off += nbytes
    take("packed", N * 2 * dp * 2); take("sqn", N * 4); take("isc", N * 4); take("mat", N * lds * 4); take("cred", N * ldcr * 4)
    take("cred2", N * ldcr * 4); take("rank_d", N * k1 * 4); take("rank_i", N * k1 * 4)
    take("rfull", N * k1 * 4); take("rfull_cnt", N * 4); take("rhalf", N * wh * 4); take("rhalf_cnt", N * 4)
    take("v1cols", N * cap1 * 4); take("v1vals", N * cap1 * 4); take("v1cnt", N * 4)
    return o, lds, ldcr, cap1

def view(t, off, n, dt):
    return t[off:off + n * 4].view(dt)

def run(N, D, k1, k2, seed):
    Q = N // 6; G = N - Q
    d = synth.market_like(Q=Q, G=G, D=D, n_ids=max(N // 25, 8), seed=seed, noise="hard")
    merged = torch.from_numpy(np.concatenate([d["gf"], d["qf"]], 0)).cuda()
    lab = [torch.from_numpy(d[k].astype(np.int32)).cuda() for k in ("ql", "qc", "gl", "gc")]
    o, lds, ldcr, cap1 = layout(N, D, k1, k2)
    snap = {}
    for mode in ("0", "2"):
        os.environ["REIDB200_TRI"] = mode
        ops.rerank_eval(merged, G, *lab, k1, k2, 0.3, want_dist=True, per_query=True)
        torch.cuda.synchronize()
        t = last[0]
        snap[mode] = dict(mat=view(t, o["mat"], N * lds, torch.float32).view(N, lds)[:, :N].clone(),
                          cred=view(t, o["cred"], N * ldcr, torch.float32).view(N, ldcr).clone(),
                          rank_i=view(t, o["rank_i"], N * k1, torch.int32).view(N, k1).clone(),
                          rank_d=view(t, o["rank_d"], N * k1, torch.float32).view(N, k1).clone(),
                          v1cols=view(t, o["v1cols"], N * cap1, torch.int32).view(N, cap1).clone(),
                          v1vals=view(t, o["v1vals"], N * cap1, torch.float32).view(N, cap1).clone(),
                          v1cnt=view(t, o["v1cnt"], N, torch.int32).clone())
    os.environ.pop("REIDB200_TRI")
    a, b = snap["0"], snap["2"]
    print("== N=%d D=%d k1=%d k2=%d" % (N, D, k1, k2))
    idx = torch.arange(N, device="cuda")
    stored = (idx[:, None] >> 8) >= (idx[None, :] >> 8)
    full_b = torch.where(stored, b["mat"], b["mat"].T)
    dm = (full_b != a["mat"])
    print("matrix: differing elements", int(dm.sum()), "of", N * N, " nan in tri view", int(torch.isnan(full_b).sum()))
    if int(dm.sum()):
        nz = dm.nonzero()
        blk = torch.stack([nz[:, 0] >> 8, nz[:, 1] >> 8], 1)
        ub, cnt = torch.unique(blk, dim=0, return_counts=True)
        print("   by (row block, col block):", [(int(x[0]), int(x[1]), int(c)) for x, c in zip(ub[:40], cnt[:40])])
        for r, c in nz[:8].cpu().numpy():
            print("   (%d,%d): tri %.9g full %.9g  sym-full %.9g" % (r, c, float(full_b[r, c]), float(a["mat"][r, c]), float(a["mat"][c, r])))
    asym = (a["mat"] != a["mat"].T)
    nz = asym.nonzero()
    blk = torch.stack([nz[:, 0] >> 8, nz[:, 1] >> 8], 1) if len(nz) else None
    print("full-mode matrix asymmetric elements:", int(asym.sum()), "" if blk is None else "off-diagonal-block ones: %d" % int((blk[:, 0] != blk[:, 1]).sum()))
    nc = (N + 31) // 32
    print("cred differs:", int((a["cred"][:, :nc] != b["cred"][:, :nc]).sum()), " rank_i differs:", int((a["rank_i"] != b["rank_i"]).sum()),
          " rank_d differs:", int((a["rank_d"] != b["rank_d"]).sum()), " v1cnt differs:", int((a["v1cnt"] != b["v1cnt"]).sum()))
    m = torch.arange(cap1, device="cuda")[None, :] < a["v1cnt"][:, None]
    print("v1cols differs:", int(((a["v1cols"] != b["v1cols"]) & m).sum()), " v1vals differs:", int(((a["v1vals"] != b["v1vals"]) & m).sum()),
          " rows:", torch.unique(((a["v1vals"] != b["v1vals"]) & m).nonzero()[:, 0])[:20].cpu().numpy())

for c in [(2305, 48, 16, 1, 2306), (2305, 64, 16, 1, 2306), (3000, 128, 20, 6, 3001), (2050, 48, 16, 1, 7)]:
    try:
        run(*c)
    except Exception as e:
        import traceback; traceback.print_exc()
