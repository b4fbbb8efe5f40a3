"""CPU: the oracle (oracle/reid_oracle.py) against the golden vectors that oracle/make_golden.py
produced by running the reference's own code (reid/evaluate.py, reid/faiss_utils.py,
reid/losses/utils.py, boxmot matching / linear_assignment / iou_matching / KF adapter,
modification_deepsort feature_extractor + iou_matching)."""
import numpy as np
import pytest

from oracle import reid_oracle as O


def test_dist(golden):
    g = golden("dist")
    e = O.euclidean_dist(g["x"], g["y"])
    mask = g["euclid"] > 1e-2
    np.testing.assert_allclose(e[mask], g["euclid"][mask], rtol=1e-5)
    np.testing.assert_allclose(O.cosine_dist(g["x"], g["y"]), g["cosine"], atol=2e-7)


def test_eval(golden):
    g = golden("eval")
    cmc, ap = O.evaluate_all(g["qf"], g["ql"], g["qc"], g["gf"], g["gl"], g["gc"])
    assert np.array_equal(cmc, g["cmc"])
    assert abs(ap - float(g["ap"])) < 1e-12
    assert (g["first_hit"] == -1).any(), "fixture must exercise the no-good-match sentinel"


def test_eval_from_dist_matches_score_path(golden):
    g = golden("eval")
    dist = -(g["qf"] @ g["gf"].T)
    cmc, ap = O.evaluate_all_from_dist(dist, g["ql"], g["qc"], g["gl"], g["gc"])
    np.testing.assert_allclose(cmc, g["cmc"], atol=1e-7)
    assert abs(ap - float(g["ap"])) < 1e-9


def test_jaccard(golden):
    g = golden("jaccard")
    _, rank = O.knn_l2(g["feats"], g["feats"], 20)
    assert np.array_equal(rank, g["initial_rank"])
    np.testing.assert_allclose(O.compute_jaccard_distance(g["feats"], 20, 6), g["J_k20_6"], atol=2e-6)
    np.testing.assert_allclose(O.compute_jaccard_distance(g["feats"], 10, 1), g["J_k10_1"], atol=2e-6)
    J16 = O.compute_jaccard_distance(g["feats"], 20, 6, use_float16=True)
    assert J16.dtype == np.float16
    np.testing.assert_allclose(J16.astype(np.float32), g["J_k20_6_f16"].astype(np.float32), atol=2e-3)


def test_jaccard_sparse_rows_equal_dense(golden):
    """The sparse restatement (what the CUDA kernels mirror) gives the dense reference result."""
    g = golden("jaccard")
    feats, rank = g["feats"], g["initial_rank"].astype(np.int64)
    rows = O.jaccard_sparse_rows(feats, rank, 20, 6)
    N = feats.shape[0]
    V = np.zeros((N, N), np.float32)
    for i, (c, v) in enumerate(rows):
        V[i, c] = v
    m = np.minimum(V[:, None, :], V[None, :, :]).sum(-1)
    J = np.maximum(1 - m / (2 - m), 0)
    np.testing.assert_allclose(J, g["J_k20_6"], atol=3e-6)


def test_preprocess(golden):
    g = golden("preprocess")
    img, xyxy = g["img"], g["xyxy"]
    H, W = img.shape[:2]
    assert np.array_equal(np.array([O.xywh_to_xyxy(b, W, H) for b in g["boxes_xywh"]], np.int32), xyxy)
    crops = [img[y1:y2, x1:x2] for x1, y1, x2, y2 in xyxy]
    cv = O.preprocess_deepsort(crops)
    assert np.abs(cv[g["cv_full_idx"]] - g["cv_full"]).max() <= 4.8e-7
    assert np.abs(cv[:, :, ::3, ::3] - g["cv_sub"]).max() <= 4.8e-7
    for c, want in zip(crops, g["pil_u8"]):
        assert np.array_equal(O.resize_pil_bilinear_u8(c, 256, 128), want)
    assert np.abs(O.preprocess_boxmot(crops[:1])[0] - g["pil0"]).max() <= 1e-6


def test_assoc_misc(golden):
    g = golden("assoc_misc")
    np.testing.assert_allclose(O.iou(g["bbox"], g["cands"]), g["iou"], rtol=1e-14)
    np.testing.assert_allclose(O.diou(g["bbox"], g["cands"]), g["diou"], rtol=1e-14)
    np.testing.assert_allclose(O.cosine_distance(g["a"], g["b"]), g["cosine"], atol=2e-7)
    np.testing.assert_allclose(O.pdist(g["a"], g["b"]), g["pdist"], atol=1e-5)
    np.testing.assert_allclose(O.nn_cosine_distance(g["a"], g["b"]), g["nn_cosine"], atol=2e-7)
    assert O.pdist(np.zeros((0, 4)), g["b"][:, :4]).shape == (0, 5)


def _frames(g):
    for f in range(int(g["n_frames"])):
        yield {k[len("f%d_" % f):]: g[k] for k in g.files if k.startswith("f%d_" % f)}


def test_assoc_sequence(golden):
    """Replay of a reference StrongSORT Tracker run: cost matrices and the assignments."""
    g = golden("assoc_seq")
    max_dist, max_iou = float(g["cfg_max_dist"]), float(g["cfg_max_iou_dist"])
    lam = float(g["cfg_mc_lambda"])
    n = 0
    for fr in _frames(g):
        m, ut, ud = O.match_frame(fr, max_dist, max_iou, lam)
        assert np.array_equal(np.array(sorted(m), np.int64).reshape(-1, 2), fr["matches"])
        assert sorted(ut) == fr["unmatched_tracks"].tolist()
        assert sorted(ud) == fr["unmatched_dets"].tolist()
        n += len(m)
    assert n > 20


def test_boxpairs_n2(golden):
    """N2: the oracle's pairwise box costs against the reference's iou.py / matching.py outputs (bit-exact)."""
    g = golden("boxpairs")
    for tag, dt in (("64", np.float64), ("32", np.float32)):
        a, b = g["a64"].astype(dt), g["b64"].astype(dt)
        for mode in ("iou", "giou", "diou", "ciou"):
            with np.errstate(invalid="ignore"):
                got = O.box_pairs(a, b, mode)
            assert got.dtype == dt and np.array_equal(got, g["%s_%s" % (mode, tag)], equal_nan=True), (mode, tag)
    a32, b32 = g["a64"].astype(np.float32), g["b64"].astype(np.float32)
    assert np.array_equal(O.bbox_ious(a32, b32), g["plus_one"])
    assert np.array_equal(O.iou_distance(a32, b32), g["iou_distance"])
    assert np.array_equal(O.fuse_score(g["iou_distance"], g["scores"]), g["fuse_score"])
    assert np.array_equal(O.fuse_iou(g["embedding"], a32, b32), g["fuse_iou64"])
    assert np.array_equal(O.fuse_iou(g["embedding"].astype(np.float32), a32, b32), g["fuse_iou32"])
    np.testing.assert_allclose(O.embedding_distance(g["feat_t"], g["feat_d"]), g["embedding"], atol=1e-12)
    assert O.iou_distance(np.zeros((0, 4)), b32).shape == (0, 31)


def test_ocsort_cost_n2(golden):
    g = golden("ocsort")
    iou, ang = O.ocsort_cost(g["det"], g["trk"], g["vel"], g["prev"], float(g["vdc"]))
    np.testing.assert_allclose(-(iou + ang), g["final_cost"], rtol=0, atol=1e-15)


def test_kalman_n3(golden):
    """N3: the oracle's Kalman steps replay the reference adapter's recorded sequence."""
    g = golden("kalman")
    T = g["z0"].shape[0]
    st = [list(O.kf_initiate(g["z0"][t])) for t in range(T)]
    for f in range(g["matched"].shape[0]):
        for t in range(T):
            st[t][0], st[t][1] = O.kf_predict(st[t][0], st[t][1])
            if g["matched"][f, t]:
                st[t] = list(O.kf_update(st[t][0], st[t][1], g["meas"][f, t], g["conf"][f, t]))
        np.testing.assert_allclose(np.stack([s[0] for s in st]), g["x"][f], rtol=1e-10, atol=1e-8)
        np.testing.assert_allclose(np.stack([s[1] for s in st]), g["P"][f], rtol=1e-8, atol=1e-8)
        np.testing.assert_allclose(np.stack([s[2] for s in st]), g["SI"][f], rtol=1e-8, atol=1e-12)


def test_posttricks_n4(golden):
    """N4: the oracle's flip-TTA merge, camera-bias removal, clamp + attribute add, DBSCAN and tracklet smoothing replay
    what the reference's own functions (and its sklearn dependency) produced."""
    g = golden("posttricks")
    np.testing.assert_allclose(O.tta_merge_normalize(g["e1"], g["e2"]), g["merged"], rtol=0, atol=2e-7)
    np.testing.assert_allclose(O.diminish_camera_bias(g["merged"], g["cams"]), g["unbiased"], rtol=0, atol=2e-5)
    for tag, other in (("plain", None), ("attr", g["attr_dist"]), ("tight", None)):
        d = O.clamp_add(g["J"], other)
        lab, core = O.dbscan_precomputed(d, float(g["eps_" + tag]), int(g["ms_" + tag]))
        assert np.array_equal(lab, g["lab_" + tag]) and np.array_equal(core, g["core_" + tag])
    valid = g["lab_plain"] != -1
    np.testing.assert_allclose(O.smooth_tracklets(g["unbiased"], g["merged_seqs"], valid), g["smooth"], rtol=0, atol=2e-7)


def test_hard_eval_fixture_is_a_curve_and_oracle_matches(golden):
    """eval.npz is generated on hard data: first hits spread over ranks, many AP values; the oracle reproduces the reference."""
    g = golden("eval")
    assert g["first_hit"].max() >= 5 and len(np.unique(g["first_hit"])) >= 5 and g["cmc"][0] < 0.9 < g["cmc"][-1]
    cmc, ap = O.evaluate_all(g["qf"], g["ql"], g["qc"], g["gf"], g["gl"], g["gc"])
    assert np.array_equal(cmc, g["cmc"]) and abs(ap - float(g["ap"])) < 1e-12
    for i in range(0, len(g["ql"]), 7):
        a, c = O.evaluate(g["qf"][i], g["ql"][i], g["qc"][i], g["gf"], g["gl"], g["gc"])
        f = -1 if c[0] == -1 else int(np.argmax(c))
        assert f == g["first_hit"][i] and abs(a - g["ap_per_query"][i]) < 1e-12


@pytest.mark.parametrize("tag,k1,k2", [("k20_6", 20, 6), ("k12_3", 12, 3)])
def test_rerank_eval_fixture_vs_oracle_and_tie_helper(golden, tag, k1, k2):
    """rerank_eval.npz (reference Jaccard + blend + compute_mAP): the oracle's restatement agrees, and the tie-aware
    interval helper of the GPU tests collapses to the reference's AP / first hit when the tolerance is zero."""
    import _ties as T

    g = golden("rerank_eval")
    dist = g["dist_" + tag]
    od = O.re_ranking_from_feats(g["qf"], g["gf"], k1, k2, float(g["lambda_value"]))
    np.testing.assert_allclose(od, dist, rtol=0, atol=2e-6)
    ap_q, first = O.evaluate_per_query_from_dist(dist, g["ql"], g["qc"], g["gl"], g["gc"])
    assert np.array_equal(first, g["first_" + tag]) and np.abs(ap_q - g["ap_q_" + tag]).max() < 1e-12
    assert g["first_" + tag].max() >= 3
    loose, width = T.check_per_query(dist, g["ql"], g["qc"], g["gl"], g["gc"], g["ap_q_" + tag], g["first_" + tag], rel_tol=0.0)
    assert width < 1e-9  # only exact ties (none expected) widen an interval at zero tolerance
    loose2, width2 = T.check_per_query(dist, g["ql"], g["qc"], g["gl"], g["gc"], g["ap_q_" + tag], g["first_" + tag], abs_tol=1e-3)
    assert loose2 > loose and width2 > width  # a coarse tolerance does open intervals


def test_market_full_fixture_present_and_consistent(golden):
    """The full-size configs[0] / configs[1] fixture (oracle/make_golden_fullsize.py): shapes, ranges, and the hard-data
    property the GPU test relies on (first hits beyond rank 0, CMC a curve)."""
    g = golden("market_full")
    assert g["rr_ap_q"].shape == (3368,) and g["rr_cmc"].shape == (15913,) and g["dist_vals"].shape[1] == 15913
    assert g["plain_first"].max() >= 10 and g["rr_first"].max() >= 3 and g["plain_cmc"][0] < 0.95
    assert 0.5 < float(g["rr_map"]) < 0.99 and float(g["plain_map"]) < float(g["rr_map"])
    assert abs(g["rr_ap_q"].sum() / 3368 - float(g["rr_map"])) < 1e-12
    assert np.all(np.diff(g["rr_cmc"]) >= 0) and np.all(np.diff(g["rank_dist"], axis=1) >= 0)
    assert np.array_equal(g["rank_vals"][:, 0], g["rank_rows"])  # self first


def test_lsap_restatement_equals_scipy_on_tie_heavy_matrices():
    """The restatement of scipy's rectangular_lsap.cpp (the model of the device solver's exact mode) returns scipy's own
    assignment, ties included: gated matrices (every entry above the gate replaced by one constant, as min_cost_matching
    does), coarse integer-like costs, both orientations, degenerate shapes."""
    from scipy.optimize import linear_sum_assignment

    rng = np.random.default_rng(0)
    for trial in range(250):
        T, M = int(rng.integers(1, 36)), int(rng.integers(1, 36))
        c = rng.random((T, M))
        thr = float(rng.choice([0.05, 0.2, 0.5, 2.0]))
        c[c > thr] = thr + 1e-5
        if trial % 3 == 0:
            c = np.round(c, 1)
        if trial % 7 == 0:
            c[:] = thr + 1e-5  # nothing admissible at all: the pairing is pure tie-breaking
        r0, c0 = linear_sum_assignment(c)
        r1, c1 = O.lsap_scipy_order(c)
        assert np.array_equal(r0, r1) and np.array_equal(c0, c1), (trial, T, M)
        a = O.min_cost_matching_lists(c, thr, list(range(10, 10 + T)), list(range(100, 100 + M)))
        b = O.min_cost_matching_lists(c, thr, list(range(10, 10 + T)), list(range(100, 100 + M)), solver=O.lsap_scipy_order)
        assert a == b
