"""BASELINE.json's headline configuration (Market-1501-shaped: 3368 queries x 15913 gallery rows, 512-d, k1 = 20, k2 = 6,
lambda = 0.3) at FULL size, where the oracle takes minutes: size-independent properties of the result plus sampled rows
checked against float64 arithmetic on the host."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

Q, G, D, IDS = 3368, 15913, 512, 751


@pytest.fixture(scope="module")
def market(rb):
    from reidb200 import synth

    d = synth.market_like(Q=Q, G=G, D=D, n_ids=IDS, seed=3)
    merged = torch.from_numpy(np.concatenate([d["gf"], d["qf"]], 0)).cuda()
    lab = [torch.from_numpy(d[k].astype(np.int32)).cuda() for k in ("ql", "qc", "gl", "gc")]
    return d, merged, lab


def test_initial_ranking_sampled_rows(rb, market):
    """k-NN lists of sampled rows against float64 distances on the host (ties aside: equal distances at the same rank)."""
    from reidb200 import ops

    d, merged, _ = market
    x = merged.cpu().numpy().astype(np.float64)
    _, rank = ops.jaccard(merged, 20, 6, want_rank=True)
    rank = rank.cpu().numpy()
    rng = np.random.default_rng(0)
    for r in rng.integers(0, Q + G, 48):
        dist = ((x - x[r]) ** 2).sum(1)
        want = np.argsort(dist, kind="stable")[:20]
        np.testing.assert_allclose(dist[rank[r]], dist[want], rtol=0, atol=3e-6)
        assert rank[r, 0] == r and len(set(rank[r].tolist())) == 20


def test_jaccard_matrix_properties(rb, market):
    """The full 19281 x 19281 Jaccard matrix: exactly symmetric (the reference's is), values in [0, 1], zero diagonal up to
    rounding, and every row's nearest entries are its k-reciprocal neighbours (distance < 1 only where sets overlap)."""
    from reidb200 import ops

    _, merged, _ = market
    J = ops.jaccard(merged, 20, 6)
    N = Q + G
    assert J.shape == (N, N)
    assert bool(torch.equal(J, J.t()))
    assert float(J.min()) >= 0.0 and float(J.max()) <= 1.0
    assert float(J.diagonal().abs().max()) <= 1e-6
    frac = float((J < 1).float().mean())
    assert 1e-5 < frac < 0.2   # sparse overlap structure: a few hundred related rows out of 19281


def test_fused_graph_and_materialised_evaluations_agree(rb, market):
    """The three ways to run the headline step give bit-identical CMC / mAP: fused kernel, its CUDA-graph replay, and
    jaccard + eval_cmc_map through the materialised Q x G matrix; the matrix block equals the full matrix's block."""
    from reidb200 import ops

    _, merged, lab = market
    cmc1, ap1, dist = ops.rerank_eval(merged, G, *lab, 20, 6, 0.3, want_dist=True)
    d2 = ops.jaccard(merged, 20, 6, False, 0.3, rows=(G, Q), cols=(0, G))
    cmc2, ap2 = ops.eval_cmc_map(d2, *lab, descending=False)
    assert torch.equal(dist, d2) and torch.equal(cmc1, cmc2) and torch.equal(ap1, ap2)
    plan = ops.RerankEvalPlan(G, Q, D, 20, 6, 0.3)
    plan.merged.copy_(merged)
    for dst, src in zip((plan.ql, plan.qc, plan.gl, plan.gc), lab):
        dst.copy_(src)
    cmc3, ap3 = plan.run()
    assert torch.equal(cmc3, cmc1) and torch.equal(ap3, ap1)
    ap = float(ap1[0])
    cmc = cmc1.cpu().numpy()
    assert 0.0 < ap <= 1.0 and np.all(np.diff(cmc) >= 0) and cmc[-1] <= 1.0 + 1e-6  # CMC is a cumulative curve
    # lambda = 0 block of the re-ranked distance is the Jaccard block itself
    J = ops.jaccard(merged, 20, 6)
    d0 = ops.jaccard(merged, 20, 6, False, 0.0, rows=(G, Q), cols=(0, G))
    assert torch.equal(d0, J[G:, :G])


def test_gallery_permutation_leaves_the_evaluation_unchanged(rb, market):
    """Reordering the gallery rows (features, labels and cameras together) cannot change CMC / mAP: the sparse
    accumulation is order independent (fixed-point atomics) and equal distances are the only order-dependent choice."""
    from reidb200 import ops

    _, merged, lab = market
    cmc, ap = ops.rerank_eval(merged, G, *lab, 20, 6, 0.3)
    perm = torch.from_numpy(np.random.default_rng(1).permutation(G)).cuda()
    merged_p = torch.cat([merged[:G][perm], merged[G:]], 0)
    cmc_p, ap_p = ops.rerank_eval(merged_p, G, lab[0], lab[1], lab[2][perm], lab[3][perm], 20, 6, 0.3)
    assert abs(float(ap_p[0]) - float(ap[0])) <= 1e-5  # (a pair's product may round differently on the other side of the diagonal)
    assert float((cmc_p - cmc).abs().max()) <= 2.0 / Q


def test_dbscan_invariants_at_full_size(rb, market):
    """DBSCAN on the full Jaccard matrix: sampled rows checked against the definition on the host (core = enough
    neighbours; a core point's neighbours carry its label or, if they are border points adjacent to several clusters, a
    lower one; noise has no core neighbour); cluster numbers are 0..n-1 in order of their lowest core index."""
    from reidb200 import ops

    _, merged, _ = market
    J = ops.jaccard(merged, 20, 6)
    eps, ms = 0.5, 7
    labels, ncl, core = ops.dbscan_precomputed(J, eps, ms)
    lab, cr, n = labels.cpu().numpy(), core.cpu().numpy().astype(bool), int(ncl.item())
    assert n > 100 and set(np.unique(lab[lab >= 0]).tolist()) == set(range(n))
    first_core = np.full(n, len(lab), np.int64)
    idx = np.nonzero(cr)[0]
    np.minimum.at(first_core, lab[idx], idx)
    assert np.all(np.diff(first_core) > 0)  # numbered by lowest core index (sklearn's search order)
    deg = (J <= eps).sum(1).cpu().numpy()
    assert np.array_equal(cr, deg >= ms)
    rng = np.random.default_rng(2)
    for r in rng.integers(0, Q + G, 300):
        nb = np.nonzero(J[r].cpu().numpy() <= np.float32(eps))[0]
        if cr[r]:
            assert np.all(lab[nb][cr[nb]] == lab[r])          # core neighbours are in the same cluster
            assert np.all(lab[nb] >= 0) and np.all(lab[nb][~cr[nb]] <= lab[r])
        elif cr[nb].any():
            assert lab[r] == lab[nb][cr[nb]].min()             # border: lowest-numbered neighbouring cluster
        else:
            assert lab[r] == -1


# ------------------------------------------------------------------------------------------------------------------
# BASELINE's headline configuration against the REFERENCE's outputs on HARD data (tests/golden/market_full.npz, written by
# oracle/make_golden_fullsize.py from reid/evaluate.py, reid/faiss_utils.py and reid/losses/utils.py imported in place).
@pytest.fixture(scope="module")
def hard_market(rb, golden):
    import hashlib

    from reidb200 import synth

    g = golden("market_full")
    d = synth.market_like(Q=Q, G=G, D=D, n_ids=IDS, seed=int(g["seed"]), noise=float(g["noise"]))
    h = hashlib.sha256()
    for k in ("qf", "ql", "qc", "gf", "gl", "gc"):
        h.update(np.ascontiguousarray(d[k]).tobytes())
    same_bits = h.hexdigest() == str(g["digest"])
    # (a host whose numpy rounds a row norm differently regenerates the features to within 1 ulp: every check below is a
    #  tolerance / tie-aware comparison, so it still applies; labels and cameras are integer draws and must be identical)
    merged = torch.from_numpy(np.concatenate([d["gf"], d["qf"]], 0)).cuda()
    lab = [torch.from_numpy(d[k].astype(np.int32)).cuda() for k in ("ql", "qc", "gl", "gc")]
    return g, d, merged, lab, same_bits


def test_headline_plain_evaluation_vs_reference_fixture(rb, hard_market):
    """configs[0]: evaluate_all at 3368 x 15913 x 512 on hard data = the reference's CMC / mAP at 1e-6 and the reference's
    per-query AP / first-hit rank for all 3368 queries."""
    import _ties as T
    from reidb200 import evaluate as E, ops

    g, d, merged, lab, _ = hard_market
    cmc, ap = E.evaluate_all(*[d[k] for k in ("qf", "ql", "qc", "gf", "gl", "gc")])
    assert np.abs(cmc.numpy() - g["plain_cmc"]).max() <= 1e-6 and abs(ap - float(g["plain_map"])) <= 1e-6
    score = ops.distmat(merged[G:], merged[:G], "dot")
    _, _, ap_q, first = ops.eval_cmc_map(score, *lab, descending=True, per_query=True)
    loose, width = T.check_per_query(-score.cpu().numpy(), d["ql"], d["qc"], d["gl"], d["gc"], g["plain_ap_q"], g["plain_first"],
                                     ap_q.cpu().numpy(), first.cpu().numpy(), rel_tol=0.0, abs_tol=2e-7)
    # (with ~17 good items per query spread over 15913 scores, a few per cent of the queries have some good item within
    #  2e-7 of a neighbour: those get an interval, every other query must be the reference's value exactly)
    exact = (np.abs(ap_q.cpu().numpy() - g["plain_ap_q"]) < 1e-9) & (first.cpu().numpy() == g["plain_first"])
    print("full-size plain: %d of %d queries bit-equal to the reference, %d inside a tie interval (total width %.3g)"
          % (int(exact.sum()), Q, loose, width))
    assert loose <= Q // 10 and int(g["plain_first"].max()) >= 10
    assert exact.sum() >= Q - loose


def test_headline_reranked_evaluation_vs_reference_fixture(rb, hard_market):
    """configs[1] at full size on hard data: sampled rows of the initial ranking, of the Jaccard matrix and of the re-ranked
    Q x G matrix against the reference's, every row sum of the Jaccard matrix, and CMC / mAP / per-query AP / first hit of
    all 3368 queries (ties within the distance tolerance aside)."""
    import _ties as T
    from reidb200 import ops

    g, d, merged, lab, same_bits = hard_market
    k1, k2, lam = int(g["k1"]), int(g["k2"]), float(g["lambda_value"])
    cmc, ap, dist, ap_q, first = ops.rerank_eval(merged, G, *lab, k1, k2, lam, want_dist=True, per_query=True)
    plan = ops.RerankEvalPlan(G, Q, D, k1, k2, lam)       # the benchmarked path: CUDA-graph replay of the fused step
    plan.merged.copy_(merged)
    for dst, src in zip((plan.ql, plan.qc, plan.gl, plan.gc), lab):
        dst.copy_(src)
    cmc_g, ap_g = plan.run()
    assert torch.equal(cmc_g, cmc) and torch.equal(ap_g, ap)
    del plan
    # (1) re-ranked matrix rows
    rows = g["dist_rows"].astype(np.int64)
    got = dist[torch.from_numpy(rows).cuda()].cpu().numpy()
    np.testing.assert_allclose(got, g["dist_vals"], rtol=0, atol=5e-6)
    # (2) Jaccard rows + all row sums
    J, rank = ops.jaccard(merged, k1, k2, want_rank=True)
    jr = g["J_rows"].astype(np.int64)
    np.testing.assert_allclose(J[torch.from_numpy(jr).cuda()].cpu().numpy(), g["J_vals"], rtol=0, atol=5e-6)
    np.testing.assert_allclose(J.sum(1, dtype=torch.float64).cpu().numpy(), g["J_row_sums"], rtol=2e-6)
    nnz = int((J < 1).sum())
    assert abs(nnz - int(g["J_nnz_lt1"])) <= 2e-4 * int(g["J_nnz_lt1"])  # which pairs share neighbours (k-NN ties aside)
    del J
    # (3) initial ranking: index mismatches only between neighbours at equal distance (1e-4 relative)
    rr = g["rank_rows"].astype(np.int64)
    got_rank = rank[torch.from_numpy(rr).cuda()].cpu().numpy()
    bad = got_rank != g["rank_vals"]
    assert bad.mean() < 5e-3
    if bad.any():
        x = merged.cpu().numpy().astype(np.float64)
        r, c = np.where(bad)
        dg = ((x[rr[r]] - x[got_rank[r, c]]) ** 2).sum(1)
        dw = ((x[rr[r]] - x[g["rank_vals"][r, c]]) ** 2).sum(1)
        np.testing.assert_allclose(dg, dw, rtol=1e-4, atol=2e-6)
    # (4) every query
    loose, width = T.check_per_query(dist.cpu().numpy(), d["ql"], d["qc"], d["gl"], d["gc"], g["rr_ap_q"], g["rr_first"],
                                     ap_q.cpu().numpy(), first.cpu().numpy(), rel_tol=0.0, abs_tol=1e-5)
    n_close = int((g["rr_margin"] < 2e-5).sum())
    assert loose <= n_close + 16, (loose, n_close)
    exact = (np.abs(ap_q.cpu().numpy() - g["rr_ap_q"]) < 1e-9) & (first.cpu().numpy() == g["rr_first"])
    assert exact.sum() >= Q - loose, "%d queries differ from the reference, %d have near-ties" % (Q - exact.sum(), loose)
    # (5) the reductions (north_star: CMC / mAP within 1e-6 absolute, ties within the distance tolerance aside)
    d_map = abs(float(ap[0]) - float(g["rr_map"]))
    d_cmc = np.abs(cmc.cpu().numpy() - g["rr_cmc"]).max()
    print("full-size re-ranked: |dmAP| = %.3g, max |dCMC| = %.3g, %d of %d queries bit-equal to the reference, %d inside a tie "
          "interval (total width %.3g), features bit-identical to the fixture's: %s"
          % (d_map, d_cmc, int(exact.sum()), Q, loose, width, same_bits))
    assert d_map <= 1e-6 + width / Q and d_cmc <= 1e-6 + loose / Q
    assert float(g["rr_cmc"][0]) < 0.97 and int(g["rr_first"].max()) >= 3
