"""GPU parity: distance matrices, k-NN, CMC/mAP through the C ABI vs the oracle and the golden vectors."""
import numpy as np
import pytest
import torch

from oracle import reid_oracle as O

pytestmark = pytest.mark.gpu

REL = 1e-4  # north_star: distances within 1e-4 relative


def _unit(x):
    return (x / np.linalg.norm(x, axis=1, keepdims=True)).astype(np.float32)


def test_distmat_golden(rb, golden):
    from reidb200 import losses_utils as U

    g = golden("dist")
    e = U.euclidean_dist(torch.from_numpy(g["x"]), torch.from_numpy(g["y"])).cpu().numpy()
    c = U.cosine_dist(torch.from_numpy(g["x"]), torch.from_numpy(g["y"])).cpu().numpy()
    mask = g["euclid"] > 1e-2  # the duplicate pair is sqrt of fp32 cancellation noise in the reference itself
    np.testing.assert_allclose(e[mask], g["euclid"][mask], rtol=REL)
    assert e[~mask].max() < 2e-2
    np.testing.assert_allclose(c, g["cosine"], rtol=REL, atol=5e-6)  # split-bf16 operands: ~2^-17 relative


@pytest.mark.parametrize("m,n,d", [(1, 1, 1), (37, 53, 48), (128, 256, 64), (129, 257, 65), (300, 1000, 512),
                                   (1000, 300, 2048), (513, 777, 100)])
@pytest.mark.parametrize("kind", ["euclid", "cosine_half", "dot", "sql2"])
def test_distmat_tc_vs_simt_vs_oracle(rb, m, n, d, kind):
    from reidb200 import ops

    rng = np.random.default_rng(m * 7 + n * 3 + d)
    x = rng.standard_normal((m, d)).astype(np.float32)
    y = rng.standard_normal((n, d)).astype(np.float32)
    tc = ops.distmat(x, y, kind).cpu().numpy()
    simt = ops.distmat_simt(x, y, kind).cpu().numpy()
    want = {"euclid": O.euclidean_dist, "cosine_half": O.cosine_dist, "sql2": O.sql2_dist,
            "dot": lambda a, b: a @ b.T}[kind](x, y)
    scale = np.abs(want).max() + 0.1
    # split-bf16 (hi+lo) tensor-core operands carry ~2^-17 relative error per product: a few 1e-6 of the scale,
    # 20x inside the 1e-4 relative tolerance of the north star
    assert np.abs(simt - want).max() <= 2e-5 * scale
    assert np.abs(tc - want).max() <= 2e-5 * scale, np.abs(tc - want).max()
    if kind in ("euclid", "sql2"):
        big = want > 0.1 * scale
        np.testing.assert_allclose(tc[big], want[big], rtol=REL)


def test_distmat_unit_features_ranking(rb):
    """Market-shaped unit features: rankings by the tensor-core distances equal fp64 rankings except ties."""
    from reidb200 import ops, synth

    d = synth.market_like(Q=200, G=3000, D=512, n_ids=150, seed=1)
    tc = ops.distmat(d["qf"], d["gf"], "euclid").cpu().numpy()
    ref = np.sqrt(np.maximum(2 - 2 * (d["qf"].astype(np.float64) @ d["gf"].astype(np.float64).T), 1e-12))
    np.testing.assert_allclose(tc, ref, rtol=REL)
    a, b = np.argsort(tc, 1, kind="stable")[:, :50], np.argsort(ref, 1, kind="stable")[:, :50]
    bad = a != b
    if bad.any():  # any disagreement must be a tie within tolerance
        r, c = np.where(bad)
        assert np.allclose(ref[r, a[r, c]], ref[r, b[r, c]], rtol=REL)


def test_distmat_accuracy_is_fp32_class(rb):
    """The scaled fp16 hi|lo split keeps 22 significand bits per operand, so the operand error is ~2^-22 relative; what is
    left is the tensor core's own fp32 accumulation, which truncates: every UMMA loses < 1 ulp of the RUNNING sum, i.e. a
    few 1e-7 for small dot products and up to ~3 * (d/16) * 2^-25 for |dot| ~ |x||y| (near duplicates)."""
    from reidb200 import ops

    rng = np.random.default_rng(3)
    x, y = _unit(rng.standard_normal((300, 512))), _unit(rng.standard_normal((700, 512)))
    y[:50] = _unit(x[:50] + 1e-3 * rng.standard_normal((50, 512)))  # near duplicates: dot ~ 1
    truth = x.astype(np.float64) @ y.astype(np.float64).T
    got = ops.distmat(x, y, "dot").cpu().numpy()
    err = np.abs(got - truth)
    assert err[np.abs(truth) < 0.2].max() <= 5e-7, err[np.abs(truth) < 0.2].max()
    assert err.max() <= 3 * (512 / 16) * 2.0 ** -25 * 2.5, err.max()  # 7.2e-6: truncating accumulation at |dot| ~ 1
    sx = (10.0 ** rng.uniform(-3, 3, (300, 1))).astype(np.float32)
    sy = (10.0 ** rng.uniform(-3, 3, (700, 1))).astype(np.float32)
    xs, ys = x * sx, y * sy
    truth = xs.astype(np.float64) @ ys.astype(np.float64).T
    got = ops.distmat(xs, ys, "dot").cpu().numpy()
    scale = np.linalg.norm(xs, axis=1)[:, None] * np.linalg.norm(ys, axis=1)[None, :]
    assert (np.abs(got - truth) <= (3e-7 + 7.2e-6 * np.abs(truth) / scale) * scale).all()
    z = np.zeros((3, 512), np.float32)  # all-zero rows: scale falls back to 1, distances stay finite
    assert np.array_equal(ops.distmat(z, y[:5], "dot").cpu().numpy(), np.zeros((3, 5), np.float32))


@pytest.mark.parametrize("n,d", [(700, 64), (1100, 512), (257, 100)])
def test_self_distance_symmetric_mode(rb, n, d, monkeypatch):
    """x vs x: only the upper triangle of 256x256 tiles is computed, off-diagonal tiles are stored twice (mirror through a
    transposed shared-memory box).  The result is exactly symmetric and matches the full computation / the oracle; the
    chunk minima of the mirrored blocks drive the same k-NN."""
    from reidb200 import ops

    rng = np.random.default_rng(n)
    x = _unit(rng.standard_normal((n, d)))
    x[n // 2] = x[3]
    xd = torch.from_numpy(x).cuda()
    for kind in ("sql2", "euclid", "dot"):
        got = ops.distmat(xd, xd, kind).cpu().numpy()
        blk = np.arange(n) // 256
        off = blk[:, None] != blk[None, :]  # mirrored tiles are copies; inside a diagonal tile both halves are computed
        assert np.array_equal(got[off], got.T[off]) and np.abs(got - got.T).max() <= 2e-5
        monkeypatch.setenv("REIDB200_GEMM_NOSYM", "1")
        full = ops.distmat(xd, xd, kind).cpu().numpy()
        monkeypatch.delenv("REIDB200_GEMM_NOSYM")
        np.testing.assert_allclose(got, full, rtol=REL, atol=2e-5)
        assert np.array_equal(np.triu(got), np.triu(full)) or np.abs(got - full).max() <= 2e-5
    np.testing.assert_allclose(ops.distmat(xd, xd, "sql2").cpu().numpy(), O.sql2_dist(x, x), rtol=REL, atol=2e-5)
    k = 20
    D, I = ops.knn(xd, xd, k, "sql2")
    monkeypatch.setenv("REIDB200_GEMM_NOSYM", "1")
    D2, I2 = ops.knn(xd, xd, k, "sql2")
    monkeypatch.delenv("REIDB200_GEMM_NOSYM")
    D, I, D2, I2 = D.cpu().numpy(), I.cpu().numpy(), D2.cpu().numpy(), I2.cpu().numpy()
    np.testing.assert_allclose(D, D2, rtol=REL, atol=2e-5)
    bad = I != I2
    if bad.any():
        fullm = O.sql2_dist(x, x).astype(np.float64)
        r, c = np.where(bad)
        assert np.allclose(fullm[r, I[r, c]], fullm[r, I2[r, c]], rtol=REL, atol=2e-5)


def test_distmat_errors(rb):
    from reidb200 import ops

    with pytest.raises(ValueError):
        ops.distmat(np.zeros((3, 4), np.float32), np.zeros((3, 5), np.float32))
    assert ops.distmat(np.zeros((0, 4), np.float32), np.zeros((3, 4), np.float32)).shape == (0, 3)


@pytest.mark.parametrize("nq,nb,d,k", [(5, 40, 16, 7), (260, 260, 32, 20), (64, 5000, 128, 100), (33, 70000, 64, 10),
                                       (7, 300, 32, 300)])
def test_knn_vs_oracle(rb, nq, nb, d, k):
    from reidb200 import ops

    rng = np.random.default_rng(nq + nb)
    xb = _unit(rng.standard_normal((nb, d)))
    xq = _unit(rng.standard_normal((nq, d)))
    if nb >= 50:
        xb[17] = xb[3]  # exact duplicates: tie -> lower index first
    D, I = ops.knn(xq, xb, k, "sql2")
    D, I = D.cpu().numpy(), I.cpu().numpy()
    Do, Io = O.knn_l2(xb, xq, k)
    np.testing.assert_allclose(D, Do, rtol=REL, atol=2e-5)
    bad = I != Io
    if bad.any():  # indices may differ only between candidates whose true distances tie within the tolerance
        full = O.sql2_dist(xq, xb).astype(np.float64)
        r, c = np.where(bad)
        assert np.allclose(full[r, I[r, c]], full[r, Io[r, c]], rtol=REL, atol=2e-5), "non-tie index mismatch"
    if nb >= 50 and k >= 2:  # the exact duplicate rows 3 and 17 always come out lower index first
        for r in range(nq):
            hit = np.where((I[r] == 3) | (I[r] == 17))[0]
            if len(hit) == 2:
                assert I[r, hit[0]] == 3
    assert (np.diff(D, axis=1) >= 0).all()


def test_knn_chunk_guided_overflow_redo(rb):
    """Thousands of exact ties overflow the warp kernel's candidate buffer: those rows are redone by the exact kernel and
    still come out lowest-index first; ordinary rows of the same launch are unaffected."""
    from reidb200 import ops

    rng = np.random.default_rng(12)
    nb, d, k = 8000, 64, 10
    xb = _unit(rng.standard_normal((nb, d)))
    xb[1000:4000] = xb[1000]
    xq = _unit(rng.standard_normal((6, d)))
    xq[0] = xb[1000]
    xq[3] = xb[2500]
    for metric in ("sql2", "dot"):
        D, I = ops.knn(xq, xb, k, metric)
        D, I = D.cpu().numpy(), I.cpu().numpy()
        assert np.array_equal(I[0], np.arange(1000, 1000 + k)) and np.array_equal(I[3], np.arange(1000, 1000 + k))
        full = O.sql2_dist(xq, xb) if metric == "sql2" else -(xq @ xb.T)
        Io = np.argsort(full, axis=1, kind="stable")[:, :k]
        bad = I != Io
        r, c = np.where(bad)
        assert np.allclose(full[r, I[r, c]], full[r, Io[r, c]], rtol=REL, atol=2e-5)


def test_knn_golden_rank(rb, golden):
    from reidb200 import ops

    g = golden("jaccard")
    _, I = ops.knn(g["feats"], g["feats"], 20, "sql2", idx64=False)
    I = I.cpu().numpy()
    assert I.dtype == np.int32 and np.array_equal(I[:, 0], np.arange(len(I)))  # self first
    bad = I != g["initial_rank"]
    assert bad.mean() < 0.01
    full = O.sql2_dist(g["feats"], g["feats"]).astype(np.float64)
    r, c = np.where(bad)
    assert np.allclose(full[r, I[r, c]], full[r, g["initial_rank"][r, c]], rtol=REL, atol=2e-5)


def test_topk_ties_and_fallback(rb):
    """Constant rows (all ties) overflow the candidate buffer and take the bisection fallback."""
    from reidb200 import ops

    v = np.zeros((3, 9000), np.float32)
    v[1, 100:] = 1.0
    v[2] = np.arange(9000, 0, -1)
    D, I = ops.topk_rows(torch.from_numpy(v), 25)
    D, I = D.cpu().numpy(), I.cpu().numpy()
    assert np.array_equal(I[0], np.arange(25)) and (D[0] == 0).all()
    assert np.array_equal(I[1], np.arange(25))
    assert np.array_equal(I[2], np.arange(8999, 8999 - 25, -1))
    D, I = ops.topk_rows(torch.from_numpy(v), 25, descending=True)
    assert np.array_equal(I.cpu().numpy()[2], np.arange(25))


def test_eval_golden(rb, golden):
    from reidb200 import evaluate as E

    g = golden("eval")
    t = {k: torch.from_numpy(g[k]) for k in ("qf", "ql", "qc", "gf", "gl", "gc")}
    cmc, ap = E.evaluate_all(t["qf"], t["ql"], t["qc"], t["gf"], t["gl"], t["gc"])
    assert cmc.dtype == torch.float32 and tuple(cmc.shape) == (g["gf"].shape[0],)
    np.testing.assert_allclose(cmc.numpy(), g["cmc"], atol=1e-6)  # north_star: CMC/mAP within 1e-6 absolute
    assert abs(ap - float(g["ap"])) < 1e-6
    # per query
    for i in (0, 5, int(np.argmin(g["first_hit"]))):
        a, c = E.evaluate(t["qf"][i], t["ql"][i], t["qc"][i], t["gf"], t["gl"], t["gc"])
        if g["first_hit"][i] < 0:
            assert c[0] == -1
        else:
            assert abs(a - g["ap_per_query"][i]) < 1e-9
            assert int(np.argmax(c.numpy())) == g["first_hit"][i]


def test_eval_market_shape_vs_oracle(rb):
    from reidb200 import evaluate as E, synth

    d = synth.market_like(Q=300, G=4000, D=128, n_ids=200, seed=2, orphan_frac=0.03)
    cmc, ap = E.evaluate_all(*[torch.from_numpy(d[k]) for k in ("qf", "ql", "qc", "gf", "gl", "gc")])
    ocmc, oap = O.evaluate_all(d["qf"], d["ql"], d["qc"], d["gf"], d["gl"], d["gc"])
    np.testing.assert_allclose(cmc.numpy(), ocmc, atol=1e-6)
    assert abs(ap - oap) < 1e-6


def test_compute_map_entry(rb):
    from reidb200 import evaluate as E

    index = np.array([4, 2, 7, 0, 1, 3, 5, 6])
    good, junk = np.array([7, 3]), np.array([2, 6])
    ap, cmc = E.compute_mAP(index, good, junk)
    oap, ocmc = O.compute_mAP(index, good, junk)
    assert abs(ap - oap) < 1e-12 and np.array_equal(cmc.numpy(), ocmc)
    ap, cmc = E.compute_mAP(index, np.array([], np.int64), junk)
    assert cmc[0] == -1
