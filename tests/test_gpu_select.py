"""GPU tests of the selection fused into the tcgen05 GEMM epilogue (csrc/gemm.cu epilogue_select, csrc/topk.cu
select_lists_kernel / sel_redo_kernel): per-row thresholds + candidate lists instead of a distance block.  The library picks
it for large galleries / merged sets; REIDB200_KNN_SELECT / REIDB200_RERANK_SELECT force either mode, and the two modes must
return the same neighbours (ties -> lower index) and, in the re-ranking step, the same matrices up to the rounding of the
pair values that no longer come from a matrix."""
import numpy as np
import pytest
import torch

from oracle import reid_oracle as O

pytestmark = pytest.mark.gpu


def _unit(x):
    return (x / np.linalg.norm(x, axis=1, keepdims=True)).astype(np.float32)


@pytest.mark.parametrize("nq,nb,d,k,metric", [(300, 5000, 64, 20, "sql2"), (7, 900, 48, 5, "sql2"), (700, 3000, 128, 100, "sql2"),
                                             (130, 4100, 64, 20, "dot"), (257, 257, 32, 20, "sql2"), (40, 15, 16, 20, "sql2"),
                                             (1500, 20000, 64, 1, "sql2")])
def test_knn_select_equals_materialised(rb, monkeypatch, nq, nb, d, k, metric):
    from reidb200 import ops

    rng = np.random.default_rng(nq + nb)
    xb = _unit(rng.standard_normal((nb, d)))
    xq = _unit(rng.standard_normal((nq, d)) + 0.5 * xb[rng.integers(0, nb, nq)])
    monkeypatch.setenv("REIDB200_KNN_SELECT", "0")
    D0, I0 = ops.knn(xq, xb, k, metric)
    monkeypatch.setenv("REIDB200_KNN_SELECT", "1")
    D1, I1 = ops.knn(xq, xb, k, metric, index_base=7, idx64=False)
    assert torch.equal(D0, D1)
    assert torch.equal(I0, I1.long() - 7 * (I1 >= 0).long())
    full = O.sql2_dist(xq, xb) if metric == "sql2" else -(xq @ xb.T)
    kk = min(k, nb)
    Io = np.argsort(full, axis=1, kind="stable")[:, :kk]
    got = I0.cpu().numpy()[:, :kk]
    r, c = np.where(got != Io)
    assert np.allclose(full[r, got[r, c]], full[r, Io[r, c]], rtol=1e-4, atol=2e-5)  # only ties differ from float64-free numpy
    if k > nb:
        assert (I1[:, nb:] == -1).all() and torch.isinf(D1[:, nb:]).all()


def test_knn_select_massive_ties_take_the_exact_redo(rb, monkeypatch):
    """Thousands of identical gallery rows: the candidate lists overflow (or hold more exact ties than the selection's
    buffer) and those rows are redone exactly; lowest index first, like the materialised path."""
    from reidb200 import ops

    rng = np.random.default_rng(12)
    nb, d, k = 9000, 64, 10
    xb = _unit(rng.standard_normal((nb, d)))
    xb[1000:5000] = xb[1000]
    xq = _unit(rng.standard_normal((300, d)))
    xq[0] = xb[1000]
    xq[3] = xb[2500]
    monkeypatch.setenv("REIDB200_KNN_SELECT", "1")
    for metric in ("sql2", "dot"):
        D, I = ops.knn(xq, xb, k, metric)
        I = I.cpu().numpy()
        assert np.array_equal(I[0], np.arange(1000, 1000 + k)) and np.array_equal(I[3], np.arange(1000, 1000 + k))
        full = O.sql2_dist(xq, xb) if metric == "sql2" else -(xq @ xb.T)
        Io = np.argsort(full, axis=1, kind="stable")[:, :k]
        r, c = np.where(I != Io)
        assert np.allclose(full[r, I[r, c]], full[r, Io[r, c]], rtol=1e-4, atol=2e-5)


@pytest.mark.parametrize("N,D,k1,k2", [(1000, 64, 20, 6), (3000, 128, 20, 6), (700, 32, 12, 3), (5200, 64, 20, 6)])
def test_rerank_select_mode_vs_matrix_mode(rb, monkeypatch, N, D, k1, k2):
    """The whole re-ranking step with the k-NN selected inside the GEMM epilogue (no N x N matrix; pair values from the
    rows' own top-k1 lists or direct dot products) against matrix mode and the oracle."""
    from reidb200 import ops, synth

    Q = N // 6
    G = N - Q
    d = synth.market_like(Q=Q, G=G, D=D, n_ids=max(N // 25, 8), seed=N, noise="hard")
    merged = torch.from_numpy(np.concatenate([d["gf"], d["qf"]], 0)).cuda()
    lab = [torch.from_numpy(d[k].astype(np.int32)).cuda() for k in ("ql", "qc", "gl", "gc")]
    out = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("REIDB200_RERANK_SELECT", mode)
        J, rank = ops.jaccard(merged, k1, k2, want_rank=True)
        cmc, ap, dist, ap_q, first = ops.rerank_eval(merged, G, *lab, k1, k2, 0.3, want_dist=True, per_query=True)
        d2 = ops.jaccard(merged, k1, k2, False, 0.3, rows=(G, Q), cols=(0, G))
        assert torch.equal(dist, d2)  # fused evaluation = two calls, in either mode
        cmc2, ap2 = ops.rerank_eval(merged, G, *lab, k1, k2, 0.3)
        assert torch.equal(cmc, cmc2) and torch.equal(ap, ap2)  # repeatable: the selection does not depend on timing
        out[mode] = (J, rank, cmc, ap, dist)
    J0, r0, cmc0, ap0, dist0 = out["0"]
    J1, r1, cmc1, ap1, dist1 = out["1"]
    assert torch.equal(r0, r1)                                   # the initial rankings are the same lists
    assert float((J0 - J1).abs().max()) <= 2e-6 and float((dist0 - dist1).abs().max()) <= 2e-6
    assert abs(float(ap0[0]) - float(ap1[0])) <= 2e-4 and float((cmc0 - cmc1).abs().max()) <= 3.0 / Q
    if N <= 1000:
        Jo = O.compute_jaccard_distance(merged.cpu().numpy(), k1, k2)
        np.testing.assert_allclose(J1.cpu().numpy(), Jo, atol=5e-6)
