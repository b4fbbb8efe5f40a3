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
