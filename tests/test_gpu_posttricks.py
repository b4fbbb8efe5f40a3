"""N4: the steps around compute_jaccard_distance on the device-resident matrix, against the golden vectors
of the reference's own functions / sklearn and against the oracle on larger seeded cases."""
import numpy as np
import pytest
import torch

from oracle import reid_oracle as O

pytestmark = pytest.mark.gpu


def test_tta_merge_and_clamp_add_golden(rb, golden):
    from reidb200 import ops

    g = golden("posttricks")
    out = ops.tta_merge_normalize(torch.from_numpy(g["e1"]), torch.from_numpy(g["e2"]))
    np.testing.assert_allclose(out.cpu().numpy(), g["merged"], rtol=0, atol=2e-7)  # row-norm summation order
    J = g["J"].copy()
    J[3, 5] = -0.25  # the clamp
    for other in (None, g["attr_dist"]):
        d = ops.clamp_add_(torch.from_numpy(J).cuda(), None if other is None else torch.from_numpy(other))
        assert np.array_equal(d.cpu().numpy(), O.clamp_add(J, other))  # bit-exact: one max, one add
    # ragged: odd width, padded rows (no float4 path)
    rng = np.random.default_rng(0)
    a = rng.standard_normal((13, 37)).astype(np.float32)
    b = rng.standard_normal((13, 37)).astype(np.float32)
    buf = torch.zeros((13, 41), device="cuda")
    view = buf[:, :37]
    view.copy_(torch.from_numpy(a))
    ops.clamp_add_(view, torch.from_numpy(b))
    assert np.array_equal(view.cpu().numpy(), O.clamp_add(a, b)) and float(buf[:, 37:].abs().sum()) == 0.0
    # contiguous with an odd element count: the flat float4 path and its scalar tail
    flat = ops.clamp_add_(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda())
    assert np.array_equal(flat.cpu().numpy(), O.clamp_add(a, b))
    assert np.array_equal(ops.clamp_add_(torch.from_numpy(a).cuda()).cpu().numpy(), O.clamp_add(a))


@pytest.mark.parametrize("tag", ["plain", "attr", "tight"])
def test_dbscan_golden(rb, golden, tag):
    """Labels, cluster numbering and core set of sklearn's DBSCAN, bit for bit."""
    from reidb200 import cluster, ops

    g = golden("posttricks")
    d = O.clamp_add(g["J"], g["attr_dist"] if tag == "attr" else None)
    cm = cluster.DBSCAN(eps=float(g["eps_" + tag]), min_samples=int(g["ms_" + tag]), metric="precomputed", n_jobs=-1)
    lab = cm.fit_predict(d)
    assert lab.dtype == np.int64 and np.array_equal(lab, g["lab_" + tag])
    assert np.array_equal(cm.core_sample_indices_, np.nonzero(g["core_" + tag])[0])
    labels, ncl, core = ops.dbscan_precomputed(torch.from_numpy(d).cuda(), float(g["eps_" + tag]), int(g["ms_" + tag]))
    assert int(ncl.item()) == int(g["lab_" + tag].max()) + 1
    with pytest.raises(ValueError):
        cluster.DBSCAN(metric="euclidean")


def test_dbscan_vs_oracle_on_jaccard(rb):
    """Larger seeded cases on the device Jaccard matrix itself (rows contiguous, or padded by hand: the scalar path when
    the stride is not a multiple of four); together they hold core, border and noise points."""
    from reidb200 import ops

    borders = 0
    for N, eps, ms, pad in ((700, 0.45, 4, 0), (1234, 0.55, 6, 3), (997, 0.7, 12, 0), (1500, 0.8, 25, 0)):
        rng = np.random.default_rng(N)
        cent = rng.standard_normal((N // 12, 48)).astype(np.float32)
        f = cent[rng.integers(0, len(cent), N)] + np.where(rng.random(N) < 0.3, 1.1, 0.45)[:, None] * rng.standard_normal((N, 48))
        f = (f / np.linalg.norm(f, axis=1, keepdims=True)).astype(np.float32)
        J = ops.jaccard(torch.from_numpy(f).cuda(), 20, 6)
        if pad:
            buf = torch.empty((N, N + pad), device="cuda")
            buf[:, :N] = J
            J = buf[:, :N]
        labels, ncl, core = ops.dbscan_precomputed(J, eps, ms)
        olab, ocore = O.dbscan_precomputed(J.cpu().numpy(), eps, ms)
        assert np.array_equal(core.cpu().numpy().astype(bool), ocore), (N, eps, ms)
        assert np.array_equal(labels.cpu().numpy().astype(np.int64), olab), (N, eps, ms)
        assert int(ncl.item()) == olab.max() + 1
        assert (olab == -1).sum() > 0 and olab.max() >= 1
        borders += int(((olab != -1) & ~ocore).sum())
    assert borders > 0  # the cases exercise every kind of point


def test_dbscan_border_takes_lowest_cluster(rb):
    """A hand-built matrix: two clusters of core points and border points within eps of both (-> the lower-numbered
    cluster, the first depth-first search that reaches them) or of one of them."""
    from reidb200 import ops

    n = 40
    d = np.ones((n, n), dtype=np.float32)
    np.fill_diagonal(d, 0.0)
    A, B = list(range(2, 12)), list(range(20, 30))     # two groups of mutually close points (core for min_samples = 5)
    for grp in (A, B):
        for i in grp:
            for j in grp:
                d[i, j] = 0.0 if i == j else 0.1
    for b_pt, cores in ((0, [25, 5]), (1, [27]), (35, [3]), (36, [22, 9]), (39, [])):  # border (and one noise) points
        for c in cores:
            d[b_pt, c] = d[c, b_pt] = 0.2
    labels, ncl, core = ops.dbscan_precomputed(torch.from_numpy(d).cuda(), 0.3, 5)
    olab, ocore = O.dbscan_precomputed(d, 0.3, 5)
    assert np.array_equal(labels.cpu().numpy().astype(np.int64), olab) and np.array_equal(core.cpu().numpy().astype(bool), ocore)
    assert olab[0] == 0 and olab[1] == 1 and olab[35] == 0 and olab[36] == 0 and olab[39] == -1 and int(ncl.item()) == 2
    try:
        from sklearn.cluster import DBSCAN
    except ImportError:
        return
    assert np.array_equal(DBSCAN(eps=0.3, min_samples=5, metric="precomputed").fit_predict(d), olab)


def test_dbscan_edges(rb):
    from reidb200 import ops

    # all points far apart: everything is noise; all points identical: one cluster
    far = torch.ones((50, 50), device="cuda") - torch.eye(50, device="cuda")
    labels, ncl, core = ops.dbscan_precomputed(far, 0.5, 2)
    assert int(ncl.item()) == 0 and bool((labels == -1).all()) and int(core.sum()) == 0
    labels, ncl, core = ops.dbscan_precomputed(torch.zeros((50, 50), device="cuda"), 0.5, 50)
    assert int(ncl.item()) == 1 and bool((labels == 0).all()) and int(core.sum()) == 50
    labels, ncl, core = ops.dbscan_precomputed(torch.zeros((333, 333), device="cuda"), 0.5, 5)  # overflowing lists
    assert int(ncl.item()) == 1 and bool((labels == 0).all()) and int(core.sum()) == 333
    # min_samples = 1: every point is core, singletons are clusters numbered by index
    labels, ncl, _ = ops.dbscan_precomputed(far, 0.5, 1)
    assert labels.cpu().tolist() == list(range(50)) and int(ncl.item()) == 50
    # rows with more neighbours than the kernel's per-row list holds (64): the union and border passes scan them again
    n = 260
    d = np.ones((n, n), dtype=np.float32)
    d[:130, :130] = 0.1          # 130 mutually close points: core for min_samples = 120
    d[200, :100] = d[:100, 200] = 0.2   # 101 neighbours (self included): a border point with an overflowing list
    d[196:206, 196:206] = np.minimum(d[196:206, 196:206], 0.3)  # ten close points (incl. 200): too few to be core
    np.fill_diagonal(d, 0.0)
    labels, ncl, core = ops.dbscan_precomputed(torch.from_numpy(d).cuda(), 0.5, 120)
    olab, ocore = O.dbscan_precomputed(d, 0.5, 120)
    assert np.array_equal(labels.cpu().numpy().astype(np.int64), olab) and np.array_equal(core.cpu().numpy().astype(bool), ocore)
    assert int(ncl.item()) == 1 and olab[200] == 0 and not ocore[200] and olab[150] == -1 and olab[198] == -1 and ocore[:130].all()
    labels, ncl, _ = ops.dbscan_precomputed(torch.zeros((0, 0), device="cuda"), 0.5, 3)
    assert labels.numel() == 0
    with pytest.raises(ValueError):
        ops.dbscan_precomputed(torch.zeros((4, 5), device="cuda"), 0.5, 3)


def test_smooth_tracklets_and_camera_bias_golden(rb, golden):
    from reidb200 import inference_utils as IU

    g = golden("posttricks")
    valid = g["lab_plain"] != -1
    # host tensors in, modified in place and returned (the reference's behaviour)
    e = torch.from_numpy(g["unbiased"].copy())
    out = IU.smooth_tracklets(e, torch.from_numpy(g["merged_seqs"]), valid)
    assert out is e
    np.testing.assert_allclose(e.numpy(), g["smooth"], rtol=0, atol=2e-7)  # group-mean summation order
    assert np.array_equal(e.numpy()[~valid], g["unbiased"][~valid])  # noise rows untouched
    # device tensors stay on the device
    ed = torch.from_numpy(g["unbiased"].copy()).cuda()
    IU.smooth_tracklets(ed, torch.from_numpy(g["merged_seqs"]).cuda(), torch.from_numpy(valid).cuda())
    np.testing.assert_allclose(ed.cpu().numpy(), g["smooth"], rtol=0, atol=2e-7)
    m = torch.from_numpy(g["merged"].copy())
    out = IU.diminish_camera_bias(m, torch.from_numpy(g["cams"]).float())
    assert out is m
    # float32 Gram + LU inverse on the device vs the reference's on the host: 2e-5 on unit-norm rows
    np.testing.assert_allclose(m.numpy(), g["unbiased"], rtol=0, atol=2e-5)


def test_center_columns_and_larger_bias_vs_oracle(rb):
    from reidb200 import inference_utils as IU, ops

    rng = np.random.default_rng(4)
    x = rng.standard_normal((1003, 130)).astype(np.float32) + 3.0
    c, mean = ops.center_columns(torch.from_numpy(x), want_mean=True)
    m64 = x.astype(np.float64).mean(0)
    # float32 accumulation of 1003 values around 3.0 (eight partial sums per column): 5e-6 against the float64 mean
    np.testing.assert_allclose(mean.cpu().numpy(), m64, rtol=0, atol=5e-6)
    np.testing.assert_allclose(c.cpu().numpy(), x - m64[None, :], rtol=0, atol=6e-6)
    f = rng.standard_normal((1500, 128)).astype(np.float32)
    f /= np.linalg.norm(f, axis=1, keepdims=True)
    cams = rng.integers(0, 6, 1500)
    got = IU.diminish_camera_bias(torch.from_numpy(f.copy()).cuda(), torch.from_numpy(cams).cuda()).cpu().numpy()
    np.testing.assert_allclose(got, O.diminish_camera_bias(f, cams), rtol=0, atol=2e-5)


def test_cluster_smooth_evaluate_pipeline_vs_oracle(rb):
    """image_reid_inference.py:246-315 end to end on the device against the same chain of oracle functions."""
    from reidb200 import inference_utils as IU, synth

    d = synth.market_like(Q=90, G=510, D=64, n_ids=30, seed=77)
    rng = np.random.default_rng(78)
    flip = lambda e: (e + 0.03 * rng.standard_normal(e.shape)).astype(np.float32)
    g2, q2 = flip(d["gf"]), flip(d["qf"])
    gseq, qseq = rng.integers(0, 4, 510), rng.integers(0, 4, 90)
    attr = rng.integers(1, 3, (30, 30)).astype(np.float32)
    ids = np.concatenate([d["gl"], d["ql"]]) % 30
    a = attr[ids] / np.linalg.norm(attr[ids], axis=1, keepdims=True)
    attr_dist = O.euclidean_dist(a.astype(np.float32), a.astype(np.float32))
    CMC, ap, pseudo = IU.cluster_smooth_evaluate((d["gf"], g2), (d["qf"], q2), d["gl"], d["ql"], d["gc"], d["qc"], gseq, qseq,
                                                 eps=0.6, min_samples=5, attribute_dist=attr_dist)
    merged = np.concatenate([O.tta_merge_normalize(d["gf"], g2), O.tta_merge_normalize(d["qf"], q2)], 0)
    merged = O.diminish_camera_bias(merged, np.concatenate([d["gc"], d["qc"]]))
    dists = O.clamp_add(O.compute_jaccard_distance(merged, 20, 6), attr_dist)
    olab, _ = O.dbscan_precomputed(dists, 0.6, 5)
    assert np.array_equal(pseudo, olab)
    seqs = np.concatenate([gseq, qseq]) * (olab.max() + 1) + olab
    sm = O.smooth_tracklets(merged, seqs, olab != -1)
    oCMC, oap = O.evaluate_all(sm[510:], d["ql"], d["qc"], sm[:510], d["gl"], d["gc"])
    assert np.abs(CMC.numpy() - np.asarray(oCMC)).max() <= 1.0 / 90 + 1e-6 and abs(ap - oap) <= 2e-3
