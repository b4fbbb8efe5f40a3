"""Triangle mode of the fused re-ranked evaluation (csrc/gemm.cu GemmArgs::lower, csrc/topk.cu TRI, PairSrc::tri): the
self-distance GEMM stores every element of the squared-L2 matrix once ((i >> 8) >= (c >> 8)) and the chunk-guided row
select takes the other half from two words per chunk -- or from the transposed place when a chunk holds two or more
passing elements.  Every stored value is bit-identical to the full-matrix kernel's, the selection is exact in both, so the
whole evaluation must be EQUAL, bit for bit, with the mode on (default; REIDB200_TRI=2 makes the call fail if it could not
be used) and off (REIDB200_TRI=0).  Reference: reid/faiss_utils.py:154-219 + reid/evaluate.py:78-105 through the same
oracle / golden tests as before (tests/test_gpu_rerank.py, test_gpu_hard_eval.py, test_gpu_fullsize.py run in this mode)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _both_modes(monkeypatch, merged, G, lab, k1, k2, lam=0.3):
    from reidb200 import _lib as L, ops

    # every scratch byte is 0xFF before each call: the half of the matrix that triangle mode does not store must never be
    # read (a stale copy from the previous full-matrix call on the same inputs would hide such a read)
    fresh = L.workspace

    def poisoned(nbytes, device):
        return fresh(nbytes, device).fill_(0xFF)

    monkeypatch.setattr(L, "workspace", poisoned)
    out = {}
    for mode in ("0", "2"):
        monkeypatch.setenv("REIDB200_TRI", mode)
        cmc, ap, dist, ap_q, first = ops.rerank_eval(merged, G, *lab, k1, k2, lam, want_dist=True, per_query=True)
        cmc2, ap2, ap_q2, first2 = ops.rerank_eval(merged, G, *lab, k1, k2, lam, per_query=True)  # evaluation-only row kernel
        torch.cuda.synchronize()
        assert torch.equal(cmc, cmc2) and torch.equal(ap, ap2) and torch.equal(ap_q, ap_q2) and torch.equal(first, first2)
        out[mode] = (cmc, ap, dist, ap_q, first)
    monkeypatch.delenv("REIDB200_TRI")
    monkeypatch.setattr(L, "workspace", fresh)
    for a, b in zip(out["0"], out["2"]):
        assert not bool(torch.isnan(a).any()) and torch.equal(a, b)
    return out["2"]


@pytest.mark.parametrize("N,D,k1,k2", [(3000, 128, 20, 6), (5200, 64, 20, 6), (2311, 32, 12, 3), (4113, 64, 25, 6), (2305, 48, 16, 1)])
def test_triangle_mode_equals_full_matrix_mode(rb, monkeypatch, N, D, k1, k2):
    from reidb200 import ops, synth

    Q = N // 6
    G = N - Q
    d = synth.market_like(Q=Q, G=G, D=D, n_ids=max(N // 25, 8), seed=N + 1, noise="hard")
    merged = torch.from_numpy(np.concatenate([d["gf"], d["qf"]], 0)).cuda()
    lab = [torch.from_numpy(d[k].astype(np.int32)).cuda() for k in ("ql", "qc", "gl", "gc")]
    cmc, ap, dist, ap_q, first = _both_modes(monkeypatch, merged, G, lab, k1, k2)
    # and the two-call path on the full matrix (reidb200_jaccard never uses the mode)
    d2 = ops.jaccard(merged, k1, k2, False, 0.3, rows=(G, Q), cols=(0, G))
    assert torch.equal(dist, d2)
    assert int((first >= 0).sum()) > 0 and 0.0 < float(ap[0]) <= 1.0


def test_triangle_mode_with_massive_ties_takes_the_exact_redo(rb, monkeypatch):
    """700 identical gallery rows spread over four 256-row tile blocks: their mutual distances are one value, so these rows
    hold > 512 candidates at the bound (the row select hands them to the exact CTA-per-row kernel, which reads a
    triangle-mode row through its transposed places) and every chunk among them has second == minimum (gathered)."""
    from reidb200 import synth

    N, D, k1, k2 = 3300, 32, 20, 6
    Q = 500
    G = N - Q
    d = synth.market_like(Q=Q, G=G, D=D, n_ids=40, seed=5, noise="hard")
    gf = d["gf"].copy()
    gf[300:1000] = gf[300]
    gf[1500:1530] = gf[1500]
    merged = torch.from_numpy(np.concatenate([gf, d["qf"]], 0)).cuda()
    lab = [torch.from_numpy(d[k].astype(np.int32)).cuda() for k in ("ql", "qc", "gl", "gc")]
    _both_modes(monkeypatch, merged, G, lab, k1, k2)


def test_triangle_mode_lambda_zero_and_plan_replay(rb, monkeypatch):
    """lambda = 0 (no blend block) and the CUDA-graph plan (ops.RerankEvalPlan captures whichever mode is set)."""
    from reidb200 import ops, synth

    N, D = 2600, 64
    Q = 400
    G = N - Q
    d = synth.market_like(Q=Q, G=G, D=D, n_ids=90, seed=11, noise="hard")
    merged = torch.from_numpy(np.concatenate([d["gf"], d["qf"]], 0)).cuda()
    lab = [torch.from_numpy(d[k].astype(np.int32)).cuda() for k in ("ql", "qc", "gl", "gc")]
    _both_modes(monkeypatch, merged, G, lab, 20, 6, lam=0.0)
    res = {}
    for mode in ("0", "2"):
        monkeypatch.setenv("REIDB200_TRI", mode)
        plan = ops.RerankEvalPlan(G, Q, D)
        plan.merged.copy_(merged)
        for dst, src in zip((plan.ql, plan.qc, plan.gl, plan.gc), lab):
            dst.copy_(src)
        plan.run()
        plan.run()
        torch.cuda.synchronize()
        res[mode] = (plan.cmc.clone(), plan.ap.clone())
    monkeypatch.delenv("REIDB200_TRI")
    assert torch.equal(res["0"][0], res["2"][0]) and torch.equal(res["0"][1], res["2"][1])
    cmc, ap = ops.rerank_eval(merged, G, *lab, 20, 6, 0.3)
    assert torch.equal(cmc, res["2"][0]) and torch.equal(ap, res["2"][1])
