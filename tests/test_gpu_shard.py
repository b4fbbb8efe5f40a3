"""GPU: resident packed gallery search (reidb200_knn_packed), the multi-pass merge of knn, and the sharded searcher
(world size 1 here; the NCCL path is exercised by tests/nccl_shard_check.py under torchrun and by bench.py --gpus N)."""
import os

import numpy as np
import pytest
import torch

from oracle import reid_oracle as O

pytestmark = pytest.mark.gpu
REL = 1e-4


def _unit(x):
    return (x / np.linalg.norm(x, axis=1, keepdims=True)).astype(np.float32)


def _check(D, I, xq, xb, k):
    Do, Io = O.knn_l2(xb, xq, k)
    np.testing.assert_allclose(D, Do, rtol=REL, atol=2e-5)
    bad = I != Io
    if bad.any():
        full = O.sql2_dist(xq, xb).astype(np.float64)
        r, c = np.where(bad)
        assert np.allclose(full[r, I[r, c]], full[r, Io[r, c]], rtol=REL, atol=2e-5), "non-tie index mismatch"
    assert (np.diff(D, axis=1) >= 0).all()


@pytest.mark.parametrize("nq,nb,d,k", [(9, 500, 48, 10), (130, 9000, 128, 100)])
def test_knn_packed_equals_knn(rb, nq, nb, d, k):
    from reidb200 import ops

    rng = np.random.default_rng(nb)
    xb, xq = _unit(rng.standard_normal((nb, d))), _unit(rng.standard_normal((nq, d)))
    D0, I0 = ops.knn(xq, xb, k, "sql2", index_base=1000)
    packed, sq, isc = ops.pack_rows(xb)
    D1, I1 = ops.knn_packed(xq, packed, sq, isc, nb, d, k, "sql2", index_base=1000)
    assert torch.equal(D0, D1) and torch.equal(I0, I1)  # same kernels, same operands: bit-identical
    _check(D1.cpu().numpy(), I1.cpu().numpy() - 1000, xq, xb, k)
    Dd, Id = ops.knn_packed(xq, packed, sq, isc, nb, d, k, "dot")
    ref = xq @ xb.T
    order = np.argsort(-ref, axis=1, kind="stable")[:, :k]
    np.testing.assert_allclose(Dd.cpu().numpy(), np.take_along_axis(ref, order, 1), rtol=REL, atol=2e-5)


def test_knn_multipass_merge(rb, monkeypatch):
    """Gallery wider than one distance block: per-block top-k tables merged by a final selection with explicit ids."""
    from reidb200 import ops

    rng = np.random.default_rng(5)
    nq, nb, d, k = 40, 6000, 64, 50
    xb, xq = _unit(rng.standard_normal((nb, d))), _unit(rng.standard_normal((nq, d)))
    xb[4100] = xb[7]  # duplicate rows in different blocks: lower id first after the merge
    xq[0] = xb[7]
    monkeypatch.setenv("REIDB200_KNN_BLOCK_FLOATS", str(nq * 1024))  # 1024 gallery columns per pass -> 6 passes
    D, I = ops.knn(xq, xb, k, "sql2")
    monkeypatch.delenv("REIDB200_KNN_BLOCK_FLOATS")
    D1, I1 = ops.knn(xq, xb, k, "sql2")
    D, I = D.cpu().numpy(), I.cpu().numpy()
    _check(D, I, xq, xb, k)
    assert I[0, 0] == 7 and I[0, 1] == 4100
    assert np.array_equal(I, I1.cpu().numpy()) and np.array_equal(D, D1.cpu().numpy())


def test_sharded_searcher_single_rank(rb):
    from reidb200 import shard

    rng = np.random.default_rng(9)
    nb, d, k = 5000, 96, 100
    xb, xq = _unit(rng.standard_normal((nb, d))), _unit(rng.standard_normal((33, d)))
    s = shard.ShardedSearcher(d)
    s.reserve(nb)
    for lo in range(0, nb, 1200):  # piecewise append into the preallocated packed shard
        s.append(xb[lo:lo + 1200])
    assert s.ntotal == nb
    D, I = s.search(xq, k)
    _check(D, I, xq, xb, k)
    s.reset()
    s.add_shard(xb[:10], base=70)
    D, I = s.search(xq[:3], 12)
    assert (I[:, 10:] == -1).all() and np.isinf(D[:, 10:]).all() and I[:, :10].min() == 70


def test_faiss_index_shim(rb):
    from reidb200 import faiss_utils as F

    rng = np.random.default_rng(2)
    xb, xq = _unit(rng.standard_normal((900, 32))), _unit(rng.standard_normal((11, 32)))
    idx = F.index_init_gpu(1, 32)
    idx.add(xb)
    D, I = idx.search(xq, 5)
    _check(D, I, xq, xb, 5)
    D2, I2 = F.search_index_pytorch(idx, torch.from_numpy(xq).cuda(), 5)
    assert np.array_equal(I2.cpu().numpy(), I)
    D3, I3 = F.search_raw_array_pytorch(None, torch.from_numpy(xb).cuda(), torch.from_numpy(xq).cuda(), 5)
    assert I3.dtype == torch.int32 and np.array_equal(I3.cpu().numpy(), I)


def test_sharded_evaluate_all_single_rank(rb):
    """Query-sharded evaluate_all at world size 1 == the single call (bit for bit) and the oracle (CMC/mAP 1e-6)."""
    from reidb200 import evaluate, shard, synth

    d = synth.market_like(Q=150, G=700, D=64, seed=4, orphan_frac=0.05)
    qf, ql, qc, gf, gl, gc = (d[k] for k in ('qf', 'ql', 'qc', 'gf', 'gl', 'gc'))
    cmc, ap = shard.evaluate_all_sharded(qf, ql, qc, gf, gl, gc)
    cmc1, ap1 = evaluate.evaluate_all_device(qf, ql, qc, gf, gl, gc)
    assert torch.equal(cmc, cmc1) and abs(float(ap) - float(ap1[0])) < 1e-14
    cmco, apo = O.evaluate_all(qf, ql, qc, gf, gl, gc)
    np.testing.assert_allclose(cmc.cpu().numpy(), cmco, atol=1e-6)
    assert abs(float(ap) - apo) < 1e-6
    CMC, ap2 = evaluate.evaluate_all_sharded(qf, ql, qc, gf, gl, gc, verbose=False)
    assert isinstance(CMC, torch.FloatTensor) and ap2 == float(ap)


@pytest.mark.parametrize("lam", [0.0, 0.3])
def test_rerank_rows_from_rank_equals_the_one_gpu_step(rb, lam):
    """The row-block entry a rank of the sharded re-ranking runs (given ranking tables, pair values from the top-k1 lists or
    direct dot products, blend from the block's own squared L2) against the one-GPU step: same ranking -> Jaccard block
    within 2e-6; with the fused evaluation the histogram / AP of the block's queries; search_option=2 in one process."""
    from reidb200 import faiss_utils as F, ops, synth

    d = synth.market_like(Q=120, G=900, D=64, n_ids=60, seed=5, noise="hard")
    G, Q = 900, 120
    N = G + Q
    merged = torch.from_numpy(np.concatenate([d["gf"], d["qf"]], 0)).cuda()
    lab = [torch.from_numpy(d[k].astype(np.int32)).cuda() for k in ("ql", "qc", "gl", "gc")]
    J, rank = ops.jaccard(merged, 20, 6, want_rank=True)
    Dk, Ik = ops.knn(merged, merged, 20, "sql2", idx64=False)
    assert torch.equal(Ik, rank)
    block = ops.rerank_rows_from_rank(merged, Ik, Dk, 20, 6, (200, 300), (0, N))
    assert float((block - J[200:500]).abs().max()) <= 2e-6
    # evaluation of a query sub-range, as a rank of shard.rerank_eval_sharded does it
    q0, nq = 30, 50
    blend = ops.distmat(merged[G + q0:G + q0 + nq], merged[:G], "sql2") if lam > 0 else None
    out, hist, ap_q, first = ops.rerank_rows_from_rank(merged, Ik, Dk, 20, 6, (G + q0, nq), (0, G), lam, blend,
                                                       labels=(lab[0][q0:q0 + nq], lab[1][q0:q0 + nq], lab[2], lab[3]))
    cmc1, ap1, dist1, ap_q1, first1 = ops.rerank_eval(merged, G, *lab, 20, 6, lam, want_dist=True, per_query=True)
    assert float((out - dist1[q0:q0 + nq]).abs().max()) <= 2e-6
    same = (first == first1[q0:q0 + nq])
    assert int(same.sum()) >= nq - 2 and float((ap_q - ap_q1[q0:q0 + nq])[same].abs().max()) <= 1e-3
    f = first[first >= 0].long()
    assert torch.equal(hist.long(), torch.bincount(f, minlength=G)[:G])
    if lam == 0.0:
        Jn = F.compute_jaccard_distance(merged, 20, 6, print_flag=False, search_option=2)
        assert Jn.shape == (N, N) and np.abs(Jn - J.cpu().numpy()).max() <= 2e-6
        with pytest.raises(ValueError):
            F.compute_jaccard_distance(merged[:10], 20, 6, print_flag=False, search_option=2)
