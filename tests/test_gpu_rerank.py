"""GPU parity: k-reciprocal Jaccard re-ranking vs the golden vectors (reference compute_jaccard_distance)."""
import numpy as np
import pytest
import torch

from oracle import reid_oracle as O

pytestmark = pytest.mark.gpu


def test_k_reciprocal_lists(rb, golden):
    from reidb200 import faiss_utils as F

    g = golden("jaccard")
    rank = g["initial_rank"].astype(np.int64)
    for i in (0, 17, 100, 255):
        assert np.array_equal(F.k_reciprocal_neigh(rank, i, 20), O.k_reciprocal_neigh(rank, i, 20))
        assert np.array_equal(F.k_reciprocal_neigh(rank, i, 10), O.k_reciprocal_neigh(rank, i, 10))


@pytest.mark.parametrize("tag,k1,k2,f16,tol", [("k20_6", 20, 6, False, 5e-6), ("k10_1", 10, 1, False, 5e-6),
                                               ("k20_6_f16", 20, 6, True, 4e-3)])
def test_jaccard_golden(rb, golden, tag, k1, k2, f16, tol):
    from reidb200 import faiss_utils as F

    g = golden("jaccard")
    J = F.compute_jaccard_distance(torch.from_numpy(g["feats"]), k1=k1, k2=k2, print_flag=False, search_option=0,
                                   use_float16=f16)
    want = g["J_" + tag]
    assert J.dtype == want.dtype and J.shape == want.shape
    np.testing.assert_allclose(J.astype(np.float32), want.astype(np.float32), atol=tol)
    assert (J >= 0).all() and (J <= 1).all()


def test_jaccard_larger_vs_oracle(rb):
    from reidb200 import faiss_utils as F, synth

    d = synth.market_like(Q=150, G=850, D=64, n_ids=60, seed=4)
    feats = np.concatenate([d["gf"], d["qf"]], 0)
    J = F.compute_jaccard_distance(torch.from_numpy(feats), print_flag=False)
    Jo = O.compute_jaccard_distance(feats)
    np.testing.assert_allclose(J, Jo, atol=5e-6)
    np.testing.assert_allclose(J, J.T, atol=2e-6)  # symmetric like the reference


def test_re_ranking_blend(rb):
    from reidb200 import faiss_utils as F, losses_utils as U, synth

    d = synth.market_like(Q=80, G=500, D=64, n_ids=40, seed=6)
    qf, gf = d["qf"], d["gf"]
    want = O.re_ranking_from_feats(qf, gf, 20, 6, 0.3)
    got = F.re_ranking_feats(qf, gf, 20, 6, 0.3)
    np.testing.assert_allclose(got, want, atol=1e-5)
    # classic signature from Euclidean distance matrices of unit features
    qg = U.euclidean_dist(qf, gf).cpu().numpy()
    qq = U.euclidean_dist(qf, qf).cpu().numpy()
    gg = U.euclidean_dist(gf, gf).cpu().numpy()
    got2 = F.re_ranking(qg, qq, gg, 20, 6, 0.3)
    # the self-distances on the diagonal are sqrt(noise) in fp32; rankings near ties may flip, so compare loosely
    assert np.mean(np.abs(got2 - want) < 1e-3) > 0.995


def test_jaccard_limits(rb):
    from reidb200 import faiss_utils as F

    with pytest.raises(ValueError):
        F.compute_jaccard_distance(torch.zeros(10, 8), k1=20, print_flag=False)


def test_reranked_eval_stream_matches_sync(rb):
    """The pipelined evaluator (copies of set i+1 under the kernels of set i) returns what the synchronous call does,
    for sets of different shapes sharing its slots, and both agree with the oracle QUERY BY QUERY on hard data: AP and
    first hit equal the oracle's unless a good and a non-good item tie within the distance tolerance (tests/_ties.py)."""
    import _ties as T
    from reidb200 import evaluate as E, ops, synth

    sets = [synth.market_like(Q=60 + 20 * s, G=400 + 50 * s, D=64, n_ids=100, seed=10 + s, noise=1.4) for s in range(5)]
    keys = ("qf", "ql", "qc", "gf", "gl", "gc")
    ev = E.RerankedEvalStream(depth=2)
    tickets, got = [], {}
    for i, d in enumerate(sets):
        tickets.append(ev.submit(*[d[k] for k in keys], 20, 6, 0.3))
        if i >= 1:
            got[i - 1] = ev.result(tickets[i - 1])
    got[len(sets) - 1] = ev.result(tickets[-1])
    with pytest.raises(ValueError):
        ev.result(tickets[0])  # its slot has been reused
    for i, d in enumerate(sets):
        cmc, ap = E.evaluate_all_reranked(*[d[k] for k in keys], 20, 6, 0.3, verbose=False)
        assert torch.equal(cmc, got[i][0]) and ap == got[i][1]
        Q, G = len(d["ql"]), len(d["gl"])
        merged = torch.from_numpy(np.concatenate([d["gf"], d["qf"]], 0)).cuda()
        cmc2, ap2, dist, ap_q, first = ops.rerank_eval(merged, G, d["ql"], d["qc"], d["gl"], d["gc"], 20, 6, 0.3, want_dist=True,
                                                       per_query=True)
        assert torch.equal(cmc2.cpu(), cmc) and float(ap2[0]) == ap
        odist = O.re_ranking_from_feats(d["qf"], d["gf"], 20, 6, 0.3)
        np.testing.assert_allclose(dist.cpu().numpy(), odist, rtol=0, atol=5e-6)
        oap_q, ofirst = O.evaluate_per_query_from_dist(odist, d["ql"], d["qc"], d["gl"], d["gc"])
        assert ofirst.max() >= 2  # hard data: the ranking matters
        loose, width = T.check_per_query(dist.cpu().numpy(), d["ql"], d["qc"], d["gl"], d["gc"], oap_q, ofirst, ap_q.cpu().numpy(),
                                         first.cpu().numpy(), rel_tol=0.0, abs_tol=1e-5)
        ocmc, oap = O.evaluate_all_from_dist(odist, d["ql"], d["qc"], d["gl"], d["gc"])
        assert abs(ap - oap) <= 1e-6 + width / Q and np.abs(cmc.numpy() - np.asarray(ocmc)).max() <= 1e-6 + loose / Q
        assert loose <= 3


@pytest.mark.parametrize("k1,k2", [(30, 10), (40, 1), (8, 4)])
def test_jaccard_other_k_vs_oracle(rb, k1, k2):
    """k1 / k2 outside the warp-kernel fast paths (more than 256 candidate slots per row, long merged lists) and a
    tiny k1 go through the CTA kernels; all of them must agree with the oracle."""
    from reidb200 import faiss_utils as F, synth

    d = synth.market_like(Q=120, G=680, D=64, n_ids=25, seed=21)
    feats = np.concatenate([d["gf"], d["qf"]], 0)
    J = F.compute_jaccard_distance(torch.from_numpy(feats), k1=k1, k2=k2, print_flag=False)
    Jo = O.compute_jaccard_distance(feats, k1=k1, k2=k2)
    np.testing.assert_allclose(J, Jo, atol=5e-6)


# ~16 good items per query (warp-level AP tail), ~70 (general tail), ~190 (more items with the query's id than the per-query
# label lists keep: the row kernel's own label scan)
@pytest.mark.parametrize("n_ids", [35, 8, 3])
@pytest.mark.parametrize("lam", [0.3, 0.0])
def test_fused_rerank_eval_equals_two_calls(rb, lam, n_ids):
    """ops.rerank_eval (evaluation fused into the re-rank kernel) gives the CMC / mAP of jaccard + eval_cmc_map bit for
    bit, and the matrix it optionally writes is the one jaccard writes."""
    from reidb200 import ops, synth

    G, Q = (700 if n_ids == 35 else 701), 130  # 701: the last group of four columns straddles the end of the row
    d = synth.market_like(Q=Q, G=G, D=64, n_ids=n_ids, seed=31, noise=0.6 if n_ids > 3 else 1.5)
    d["gl"][::17] = -1                      # junk gallery rows
    d["ql"][5] = 10 ** 6                    # a query without any good match: the -1 sentinel path
    merged = torch.from_numpy(np.concatenate([d["gf"], d["qf"]], 0)).cuda()
    lab = [torch.from_numpy(d[k].astype(np.int32)).cuda() for k in ("ql", "qc", "gl", "gc")]
    dist = ops.jaccard(merged, 20, 6, False, lam, rows=(G, Q), cols=(0, G))
    cmc2, ap2 = ops.eval_cmc_map(dist, *lab, descending=False)
    cmc1, ap1, dist1 = ops.rerank_eval(merged, G, *lab, 20, 6, lam, want_dist=True)
    assert torch.equal(dist1, dist) and torch.equal(cmc1, cmc2) and torch.equal(ap1, ap2)
    cmc0, ap0 = ops.rerank_eval(merged, G, *lab, 20, 6, lam)
    assert torch.equal(cmc0, cmc2) and torch.equal(ap0, ap2)


def test_rerank_eval_plan_replays_the_same_evaluation(rb):
    """ops.RerankEvalPlan (the evaluation captured as one CUDA graph) returns what the plain launches return, replay
    after replay with new inputs in its buffers; evaluate_all_reranked reuses the cached plan for a repeated shape."""
    from reidb200 import evaluate as E, ops, synth

    G, Q = 640, 110
    plan = ops.RerankEvalPlan(G, Q, 64, 20, 6, 0.3)
    for seed in (41, 42, 43):
        d = synth.market_like(Q=Q, G=G, D=64, n_ids=30, seed=seed)
        merged = torch.from_numpy(np.concatenate([d["gf"], d["qf"]], 0)).cuda()
        lab = [torch.from_numpy(d[k].astype(np.int32)).cuda() for k in ("ql", "qc", "gl", "gc")]
        plan.merged.copy_(merged)
        for dst, src in zip((plan.ql, plan.qc, plan.gl, plan.gc), lab):
            dst.copy_(src)
        cmc_g, ap_g = [t.clone() for t in plan.run()]
        cmc, ap = ops.rerank_eval(merged, G, *lab, 20, 6, 0.3)
        assert torch.equal(cmc_g, cmc) and torch.equal(ap_g, ap)
        keys = ("qf", "ql", "qc", "gf", "gl", "gc")
        cmc_h, ap_h = E.evaluate_all_reranked(*[d[k] for k in keys], 20, 6, 0.3, verbose=False)
        assert torch.equal(cmc_h, cmc.cpu()) and ap_h == float(ap.cpu()[0])
    assert len(E._plan) == 1
    E.release_plans()
    with pytest.raises(ValueError):
        ops.RerankEvalPlan(ops.RERANK_EVAL_MAX_G + 1, 4, 64)
