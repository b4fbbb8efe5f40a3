"""GPU parity, N2: pairwise box costs of the other boxmot trackers through the C ABI (csrc/boxpairs.cu)
against the reference's outputs (tests/golden/boxpairs.npz) and the oracle at MOT20 size."""
import numpy as np
import pytest

from oracle import reid_oracle as O

pytestmark = pytest.mark.gpu


class _Obj:
    def __init__(self, tlbr, score=1.0, feat=None):
        self.tlbr, self.score, self.curr_feat, self.smooth_feat = tlbr, score, feat, feat


@pytest.mark.parametrize("tag,dt", [("64", np.float64), ("32", np.float32)])
def test_batch_metrics_golden(rb, golden, tag, dt):
    from reidb200 import iou as I

    g = golden("boxpairs")
    a, b = g["a64"].astype(dt), g["b64"].astype(dt)
    for mode in ("iou", "giou", "diou"):
        got = I.get_asso_func(mode)(a, b)
        assert got.dtype == dt and np.array_equal(got, g["%s_%s" % (mode, tag)]), mode  # bit-exact
    got = I.ciou_batch(a, b)
    want = g["ciou_" + tag]
    assert got.dtype == dt and np.array_equal(np.isnan(got), np.isnan(want))  # identical boxes: 0/0 in the reference too
    ok = ~np.isnan(want)
    # arctan is a library function on both sides: last-bit differences, amplified by alpha = v / (1 - iou + v)
    np.testing.assert_allclose(got[ok], want[ok], rtol=0, atol=1e-6 if dt == np.float32 else 1e-14)


def test_plus_one_family_golden(rb, golden):
    from reidb200 import matching as M

    g = golden("boxpairs")
    a32, b32 = g["a64"].astype(np.float32), g["b64"].astype(np.float32)
    got = M.bbox_ious(a32, b32)
    assert got.dtype == np.float32 and np.array_equal(got, g["plus_one"])
    assert np.array_equal(M.ious(list(a32), list(b32)), g["plus_one"])
    d = M.iou_distance(list(a32), list(b32))
    assert d.dtype == np.float32 and np.array_equal(d, g["iou_distance"])
    trk = [_Obj(a32[i], 1.0, g["feat_t"][i]) for i in range(len(a32))]
    det = [_Obj(b32[i], float(g["scores"][i]), g["feat_d"][i]) for i in range(len(b32))]
    assert np.array_equal(M.iou_distance(trk, det), g["iou_distance"])
    fs = M.fuse_score(M.iou_distance(trk, det), det)
    assert fs.dtype == np.float64 and np.array_equal(fs, g["fuse_score"])
    f64 = M.fuse_iou(g["embedding"], trk, det)
    assert f64.dtype == np.float64 and np.array_equal(f64, g["fuse_iou64"])
    f32 = M.fuse_iou(g["embedding"].astype(np.float32), trk, det)
    assert f32.dtype == np.float32 and np.array_equal(f32, g["fuse_iou32"])
    emb = M.embedding_distance(trk, det)
    assert emb.dtype == np.float64 and emb.shape == g["embedding"].shape
    np.testing.assert_allclose(emb, g["embedding"], rtol=0, atol=1e-6)  # fp32-accurate tensor-core product vs float64 cdist


def test_empty_inputs(rb):
    from reidb200 import iou as I, matching as M

    z, b = np.zeros((0, 4), np.float32), np.ones((3, 4), np.float32)
    assert M.ious(z, b).shape == (0, 3) and M.iou_distance([], list(b)).shape == (0, 3)
    assert M.embedding_distance([], []).shape == (0, 0)
    assert I.iou_batch(np.zeros((0, 4)), np.ones((3, 4))).shape == (0, 3)
    assert M.fuse_score(np.zeros((0, 3)), []).shape == (0, 3)


def test_mot20_size_vs_oracle(rb):
    """300 tracks x 200 detections, every mode, against the oracle."""
    from reidb200 import iou as I, matching as M

    rng = np.random.default_rng(8)
    xy, wh = rng.uniform(0, 1800, (500, 2)), rng.uniform(8, 260, (500, 2))
    bx = np.concatenate([xy, xy + wh], 1)
    a, b = bx[:300], bx[300:]
    for mode in ("iou", "giou", "diou"):
        assert np.array_equal(I.get_asso_func(mode)(a, b), O.box_pairs(a, b, mode))
    np.testing.assert_allclose(I.ciou_batch(a, b), O.box_pairs(a, b, "ciou"), rtol=0, atol=1e-13)
    a32, b32 = a.astype(np.float32), b.astype(np.float32)
    assert np.array_equal(M.bbox_ious(a32, b32), O.bbox_ious(a32, b32))
    assert np.array_equal(M.iou_distance(list(a32), list(b32)), O.iou_distance(a32, b32))
    sc = rng.uniform(0, 1, 200)
    assert np.array_equal(M.fuse_score(O.iou_distance(a32, b32), sc), O.fuse_score(O.iou_distance(a32, b32), sc))


def test_ocsort_associate_golden(rb, golden):
    """OC-SORT associate: the fused IoU + velocity-direction cost equals the reference's final_cost (captured at its
    linear_assignment call) and the matches / unmatched sets equal the reference's."""
    from reidb200 import association as A

    g = golden("ocsort")
    iou, ang = A.associate_costs(g["det"], g["trk"], g["vel"], g["prev"], float(g["vdc"]))
    np.testing.assert_allclose(-(iou + ang), g["final_cost"], rtol=0, atol=1e-14)  # acos is a library function on both sides
    oi, oa = O.ocsort_cost(g["det"], g["trk"], g["vel"], g["prev"], float(g["vdc"]))
    assert np.array_equal(iou, oi)
    m, ud, ut = A.associate(g["det"], g["trk"], 0.3, g["vel"], g["prev"], float(g["vdc"]))
    assert np.array_equal(m, g["matches"])
    assert np.array_equal(np.asarray(ud), g["unmatched_dets"]) and np.array_equal(np.asarray(ut), g["unmatched_trks"])
    Y, X = A.speed_direction_batch(g["det"], g["prev"])
    assert np.array_equal(Y, g["Y"]) and np.array_equal(X, g["X"])
    m0, ud0, ut0 = A.associate(g["det"], np.zeros((0, 5)), 0.3, np.zeros((0, 2)), np.zeros((0, 5)), 0.2)
    assert m0.shape == (0, 2) and len(ud0) == len(g["det"])
