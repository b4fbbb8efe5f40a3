"""CPU, world_size 2 over gloo: host logic of the gallery-sharded search (partitioning, global ids, all_gather layout,
merge tie order).  The per-shard search and the merge are injected numpy stand-ins built on the oracle; the product
defaults are the CUDA kernels (tests/test_gpu_shard.py)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _oracle_local_search(xq, rows, k, metric, base):
    from oracle import reid_oracle as O

    kk = min(k, rows.shape[0])
    D = np.full((xq.shape[0], k), np.inf, np.float32)
    I = np.full((xq.shape[0], k), -1, np.int64)
    Dk, Ik = O.knn_l2(rows.numpy(), xq.numpy(), kk)
    D[:, :kk], I[:, :kk] = Dk, Ik + base
    return torch.from_numpy(D), torch.from_numpy(I)


def _oracle_merge(cand_d, cand_i, k, metric):
    d, i = cand_d.numpy(), cand_i.numpy()
    order = np.lexsort((i, d), axis=1)[:, :k]  # by distance, ties -> lower global id
    return torch.from_numpy(np.take_along_axis(d, order, 1)), torch.from_numpy(np.take_along_axis(i, order, 1))


def _worker(rank, world, port, n_gallery, out_dir):
    sys.path.insert(0, ROOT)
    import reidb200  # noqa: F401
    from reidb200 import shard

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(0)
        gallery = rng.standard_normal((n_gallery, 16)).astype(np.float32)
        xq = rng.standard_normal((7, 16)).astype(np.float32)
        if n_gallery > 5:
            gallery[5] = gallery[n_gallery - 2]  # an exact duplicate across the two shards: tie -> lower id first
            xq[0] = gallery[5]
        s = shard.ShardedSearcher(16, local_search=_oracle_local_search, merge=_oracle_merge)
        s.add(gallery)
        assert s.ntotal == n_gallery
        lo, hi = shard.shard_bounds(n_gallery, world)[rank]
        assert (s.base, s.n_local) == (lo, hi - lo)
        D, I = s.search(torch.from_numpy(xq), 4)
        np.savez(os.path.join(out_dir, "r%d.npz" % rank), D=D, I=I)
        # per-rank rows with implicit bases (exclusive prefix sum over ranks)
        s2 = shard.ShardedSearcher(16, local_search=_oracle_local_search, merge=_oracle_merge)
        s2.add_shard(gallery[lo:hi])
        assert s2.base == lo
        D2, I2 = s2.search(torch.from_numpy(xq), 4)
        assert np.array_equal(I2, I) and np.array_equal(D2, D)
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("n_gallery", [41, 3])
def test_sharded_search_world2_gloo(tmp_path, n_gallery):
    from oracle import reid_oracle as O

    mp.spawn(_worker, args=(2, _free_port(), n_gallery, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = np.load(tmp_path / "r0.npz"), np.load(tmp_path / "r1.npz")
    assert np.array_equal(r0["I"], r1["I"]) and np.array_equal(r0["D"], r1["D"])  # every rank returns the same table
    rng = np.random.default_rng(0)
    gallery = rng.standard_normal((n_gallery, 16)).astype(np.float32)
    xq = rng.standard_normal((7, 16)).astype(np.float32)
    if n_gallery > 5:
        gallery[5] = gallery[n_gallery - 2]
        xq[0] = gallery[5]
    k = 4
    Dw, Iw = O.knn_l2(gallery, xq, min(k, n_gallery))
    assert np.array_equal(r0["I"][:, : Iw.shape[1]], Iw)
    np.testing.assert_allclose(r0["D"][:, : Dw.shape[1]], Dw, rtol=1e-5, atol=1e-5)
    if n_gallery > 5:
        assert r0["I"][0, 0] == 5 and r0["I"][0, 1] == n_gallery - 2
    if n_gallery < k:  # fewer rows than k: padded like faiss
        assert (r0["I"][:, n_gallery:] == -1).all() and np.isinf(r0["D"][:, n_gallery:]).all()


def test_shard_bounds():
    sys.path.insert(0, ROOT)
    import reidb200  # noqa: F401
    from reidb200 import shard

    assert shard.shard_bounds(10, 4) == [(0, 3), (3, 6), (6, 9), (9, 10)]
    assert shard.shard_bounds(2, 4) == [(0, 1), (1, 2), (2, 2), (2, 2)]
    assert shard.shard_bounds(0, 2) == [(0, 0), (0, 0)]
    b = shard.shard_bounds(4_000_000, 8)
    assert b[0] == (0, 500_000) and b[-1] == (3_500_000, 4_000_000)


# ------------------------------------------------------------------ query-sharded evaluate_all (CMC/mAP row)
def _oracle_local_eval(qf, ql, qc, gf, gl, gc):
    from oracle import reid_oracle as O

    n = qf.shape[0]
    ap = np.zeros(n, np.float64)
    first = np.full(n, -1, np.int32)
    for i in range(n):
        ap_i, cmc_i = O.evaluate(qf[i].numpy(), int(ql[i]), int(qc[i]), gf.numpy(), gl.numpy(), gc.numpy())
        if cmc_i[0] == -1:
            continue
        ap[i], first[i] = ap_i, int(np.argmax(cmc_i))
    return torch.from_numpy(ap), torch.from_numpy(first)


def _eval_case(Q, G, seed=3):
    rng = np.random.default_rng(seed)
    gl = rng.integers(0, 9, G)
    gl[rng.random(G) < 0.15] = -1
    gc = rng.integers(0, 3, G)
    ql, qc = rng.integers(0, 9, Q), rng.integers(0, 3, Q)
    if Q > 2:
        ql[2] = 77  # a query without any valid match: skipped in the sums, still counted in the divisor
    unit = lambda a: (a / np.linalg.norm(a, axis=1, keepdims=True)).astype(np.float32)
    cent = rng.standard_normal((100, 24))
    gf = unit(cent[gl] + 0.8 * rng.standard_normal((G, 24)))
    qf = unit(cent[ql] + 0.8 * rng.standard_normal((Q, 24)))
    return qf, ql, qc, gf, gl, gc


def _eval_worker(rank, world, port, Q, out_dir):
    sys.path.insert(0, ROOT)
    import reidb200  # noqa: F401
    from reidb200 import shard

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        qf, ql, qc, gf, gl, gc = _eval_case(Q, 60)
        cmc, ap = shard.evaluate_all_sharded(qf, ql, qc, gf, gl, gc, local_eval=_oracle_local_eval)
        np.savez(os.path.join(out_dir, "e%d.npz" % rank), cmc=cmc.numpy(), ap=float(ap))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("Q", [11, 1])
def test_sharded_evaluate_all_world2_gloo(tmp_path, Q):
    from oracle import reid_oracle as O

    mp.spawn(_eval_worker, args=(2, _free_port(), Q, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = np.load(tmp_path / "e0.npz"), np.load(tmp_path / "e1.npz")
    assert np.array_equal(r0["cmc"], r1["cmc"]) and r0["ap"] == r1["ap"]  # same answer on every rank
    qf, ql, qc, gf, gl, gc = _eval_case(Q, 60)
    cmc, ap = O.evaluate_all(qf, ql, qc, gf, gl, gc)
    assert np.array_equal(r0["cmc"], cmc)  # integer counts / Q: exact
    assert abs(float(r0["ap"]) - ap) < 1e-12


# ------------------------------------------------------------------ row-sharded re-ranked evaluation
def _oracle_local_knn(x_rows, x_all, k1):
    from oracle import reid_oracle as O

    D, I = O.knn_l2(x_all.numpy(), x_rows.numpy(), k1)
    return torch.from_numpy(D.astype(np.float32)), torch.from_numpy(I.astype(np.int32))


def _oracle_local_rows(merged, rank_i, rank_d, k1, k2, row0, nrows, G, lam, ql, qc, gl, gc):
    """(first-hit histogram i64[G], AP f64[nrows]) of query rows [row0, row0 + nrows) of the merged set from the GLOBAL
    initial ranking, with the oracle's dense Jaccard stage."""
    from oracle import reid_oracle as O

    x = merged.numpy()
    J = O.compute_jaccard_distance(x, k1, k2, initial_rank=rank_i.numpy().astype(np.int64))
    q = x[row0:row0 + nrows]
    dist = (J[row0:row0 + nrows, :G] * np.float32(1.0 - lam) + O.euclidean_dist(q, x[:G]) * np.float32(lam)).astype(np.float32)
    ap_q, first = O.evaluate_per_query_from_dist(dist, ql.numpy(), qc.numpy(), gl.numpy(), gc.numpy())
    hist = np.bincount(first[first >= 0], minlength=G)[:G]
    return torch.from_numpy(hist), torch.from_numpy(ap_q)


def _rerank_worker(rank, world, port, Q, out_dir):
    sys.path.insert(0, ROOT)
    import reidb200  # noqa: F401
    from reidb200 import shard

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        qf, ql, qc, gf, gl, gc = _eval_case(Q, 61)
        merged = torch.from_numpy(np.concatenate([gf, qf], 0))
        ri, rd = shard.initial_rank_sharded(merged, 6, local_knn=_oracle_local_knn)
        cmc, ap = shard.rerank_eval_sharded(qf, ql, qc, gf, gl, gc, 6, 3, 0.3, local_knn=_oracle_local_knn, local_rows=_oracle_local_rows)
        np.savez(os.path.join(out_dir, "rr%d.npz" % rank), cmc=cmc.numpy(), ap=float(ap), ri=ri.numpy(), rd=rd.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("Q", [13, 1])
def test_sharded_rerank_eval_world2_gloo(tmp_path, Q):
    """Partitioning, the padded all-gather of the ranking tables (61 + Q rows over 2 ranks: unequal shards), the query
    shards and the all-reduce of first-hit counts + AP sum: every rank returns the single-process answer."""
    from oracle import reid_oracle as O

    mp.spawn(_rerank_worker, args=(2, _free_port(), Q, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = np.load(tmp_path / "rr0.npz"), np.load(tmp_path / "rr1.npz")
    assert np.array_equal(r0["cmc"], r1["cmc"]) and r0["ap"] == r1["ap"] and np.array_equal(r0["ri"], r1["ri"])
    qf, ql, qc, gf, gl, gc = _eval_case(Q, 61)
    merged = np.concatenate([gf, qf], 0)
    D, I = O.knn_l2(merged, merged, 6)
    assert np.array_equal(r0["ri"], I) and np.array_equal(r0["rd"], D.astype(np.float32))
    dist_qg = O.re_ranking_from_feats(qf, gf, 6, 3, 0.3)
    cmc, ap = O.evaluate_all_from_dist(dist_qg, ql, qc, gl, gc)
    assert np.array_equal(r0["cmc"], cmc) and abs(float(r0["ap"]) - ap) < 1e-12
