"""GPU parity: association costs and assignments vs a recorded run of the reference StrongSORT Tracker
(tests/golden/assoc_seq.npz) and the free functions (assoc_misc.npz)."""
import numpy as np
import pytest
import torch

from oracle import reid_oracle as O

pytestmark = pytest.mark.gpu


def _frames(g):
    for f in range(int(g["n_frames"])):
        yield {k[len("f%d_" % f):]: g[k] for k in g.files if k.startswith("f%d_" % f)}


def _metric_for(fr, budget, max_dist):
    from reidb200 import matching as MT

    metric = MT.NearestNeighborDistanceMetric("cosine", max_dist, budget)
    ids = fr["sample_ids"].tolist()
    if ids:
        targets = np.repeat(fr["sample_ids"], fr["sample_cnt"])
        metric.partial_fit(fr["sample_rows"], targets, ids)
    return metric


def test_misc_free_functions(rb, golden):
    from reidb200 import iou_matching as IM, matching as MT

    g = golden("assoc_misc")
    assert np.array_equal(IM.iou(g["bbox"], g["cands"]), g["iou"])      # fp64, bit-exact
    assert np.array_equal(IM.diou(g["bbox"], g["cands"]), g["diou"])
    np.testing.assert_allclose(MT._cosine_distance(g["a"], g["b"]), g["cosine"], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(MT._pdist(g["a"], g["b"]), g["pdist"], atol=1e-4)
    np.testing.assert_allclose(MT._nn_cosine_distance(g["a"], g["b"]), g["nn_cosine"], rtol=1e-4, atol=1e-5)
    assert MT._pdist(np.zeros((0, 4)), g["b"][:, :4]).shape == (0, 5)
    with pytest.raises(ValueError):
        MT.NearestNeighborDistanceMetric("manhattan", 0.2, 5)


def test_gallery_ring_semantics(rb):
    """partial_fit keeps the last `budget` rows per target in order, drops inactive targets (matching.py:374-389)."""
    from reidb200 import matching as MT

    rng = np.random.default_rng(0)
    metric = MT.NearestNeighborDistanceMetric("cosine", 0.2, 3)
    ref = {}
    for step in range(6):
        n = int(rng.integers(1, 6))
        feats = rng.standard_normal((n, 8)).astype(np.float32)
        targets = rng.integers(0, 3, n)
        active = [0, 1, 2] if step != 4 else [0, 2]
        metric.partial_fit(feats, targets, active)
        for f, t in zip(feats, targets):
            ref.setdefault(int(t), []).append(f / np.linalg.norm(f))
            ref[int(t)] = ref[int(t)][-3:]
        ref = {k: ref[k] for k in active if k in ref}
        assert sorted(metric.samples.keys()) == sorted(ref)
        for k in ref:
            np.testing.assert_allclose(np.stack(metric.samples[k]), np.stack(ref[k]), atol=1e-7)


@pytest.mark.parametrize("packed", [False, True])
def test_sequence_costs_and_assignments(rb, golden, monkeypatch, packed):
    from reidb200 import tracker as TR

    monkeypatch.setenv("REIDB200_ASSOC_V2" if packed else "REIDB200_ASSOC_V1", "1")
    if not packed:
        monkeypatch.delenv("REIDB200_ASSOC_V2", raising=False)

    g = golden("assoc_seq")
    max_dist, max_iou, lam, budget = (float(g["cfg_max_dist"]), float(g["cfg_max_iou_dist"]), float(g["cfg_mc_lambda"]),
                                      int(g["cfg_budget"]))
    n_match = 0
    for fr in _frames(g):
        metric = _metric_for(fr, budget, max_dist)
        T, M = len(fr["trk_id"]), len(fr["det_tlwh"])
        conf = np.where(fr["trk_confirmed"])[0]
        if T and M:
            app, raw, iou = TR.frame_costs(metric, fr["det_feat"], fr["det_tlwh"], fr["trk_id"], fr["trk_confirmed"],
                                           fr["trk_mean"], fr["trk_SI"], fr["trk_tsu"], lam, max_dist, max_iou, clamp=False,
                                           want_raw=True)
            app, raw, iou = app.cpu().numpy(), raw.cpu().numpy(), iou.cpu().numpy()
            assert np.array_equal(iou, fr["iou_cost"])  # fp64 box arithmetic: bit-exact
            if len(conf):
                np.testing.assert_allclose(raw[conf], fr["app_raw"], rtol=1e-4, atol=1e-5)
                np.testing.assert_allclose(app[conf], fr["app_gated"], rtol=1e-4, atol=1e-5)
                assert np.array_equal(app[conf] > 1e4, fr["app_gated"] > 1e4)  # identical gate decisions
        m, ut, ud = TR.match_arrays(metric, fr["det_feat"], fr["det_tlwh"], fr["trk_id"], fr["trk_confirmed"], fr["trk_mean"],
                                    fr["trk_SI"], fr["trk_tsu"], lam, max_iou, max_dist, device_lap=False)  # scipy on the host
        md, utd, udd = TR.match_arrays(metric, fr["det_feat"], fr["det_tlwh"], fr["trk_id"], fr["trk_confirmed"], fr["trk_mean"],
                                       fr["trk_SI"], fr["trk_tsu"], lam, max_iou, max_dist)  # default: exact assignment on the device
        assert (md, utd, udd) == (m, ut, ud)
        assert np.array_equal(np.array(sorted(m), np.int64).reshape(-1, 2), fr["matches"])  # bit-exact assignments
        assert sorted(ut) == fr["unmatched_tracks"].tolist()
        assert sorted(ud) == fr["unmatched_dets"].tolist()
        n_match += len(m)
    assert n_match > 20


@pytest.mark.parametrize("M,T", [(30, 100), (200, 300)])
def test_mot_shaped_frame_vs_oracle(rb, M, T):
    """C3/C4-shaped frame: fused kernel vs oracle cost matrices; identical Hungarian assignments."""
    from reidb200 import matching as MT, tracker as TR, synth
    from scipy.optimize import linear_sum_assignment

    fr = synth.mot_frame(M=M, T=T, budget=100, D=512, seed=5)
    metric = MT.NearestNeighborDistanceMetric("cosine", 0.2, 100, max_tracks=T)
    targets = np.repeat(np.arange(T), [len(s) for s in fr["samples"]])
    metric.partial_fit(np.concatenate(fr["samples"], 0), targets, list(range(T)))
    conf = np.ones(T, bool)
    app, raw, iou = TR.frame_costs(metric, fr["det_feat"], fr["det_tlwh"], np.arange(T), conf, fr["trk_mean"], fr["trk_SI"],
                                   fr["trk_tsu"], 0.995, 0.2, 0.7, clamp=True, want_raw=True)
    app, raw, iou = app.cpu().numpy(), raw.cpu().numpy(), iou.cpu().numpy()
    oraw = O.nn_distance_matrix(fr["samples"], fr["det_feat"])
    np.testing.assert_allclose(raw, oraw, rtol=1e-4, atol=1e-5)
    ogated = O.gate_cost_matrix(oraw.copy(), fr["trk_mean"][:, :4], fr["trk_SI"], O.tlwh_to_xyah_f32(fr["det_tlwh"]), 0.995)
    oapp = O.clamp_cost(ogated.copy(), 0.2)
    np.testing.assert_allclose(app, oapp, rtol=1e-4, atol=1e-5)
    oiou = O.clamp_cost(O.iou_cost([O.track_to_tlwh(m[:4]) for m in fr["trk_mean"]], fr["trk_tsu"], fr["det_tlwh"]), 0.7)
    assert np.array_equal(iou, oiou)
    r1, c1 = linear_sum_assignment(app)
    r2, c2 = linear_sum_assignment(oapp)
    keep1 = [(r, c) for r, c in zip(r1, c1) if app[r, c] <= 0.2]
    keep2 = [(r, c) for r, c in zip(r2, c2) if oapp[r, c] <= 0.2]
    assert keep1 == keep2 and len(keep1) > 0


@pytest.mark.parametrize("packed", [False, True])
@pytest.mark.parametrize("M,T,budget,D", [(5, 3, 4, 24), (300, 40, 100, 96), (70, 9, 150, 512), (16, 130, 100, 64), (200, 300, 100, 512),
                                          (33, 57, 128, 128), (9, 21, 31, 64)])
def test_tensor_core_kernel_vs_simt_vs_oracle(rb, monkeypatch, M, T, budget, D, packed):
    """tcgen05 frame kernel against the fp32 SIMT kernel and the oracle: > 256 detections (detection tiles), budget > 128
    (row tiles), feature dimension not a multiple of 64 (zero padded k-blocks), empty and unknown slots."""
    from reidb200 import ops, synth

    # packed: tracks packed into 128-row tiles by quadrants with count-aware loads (assoc_frame2_kernel, rings <= 128 rows)
    monkeypatch.setenv("REIDB200_ASSOC_V2" if packed else "REIDB200_ASSOC_V1", "1")
    if not packed:
        monkeypatch.delenv("REIDB200_ASSOC_V2", raising=False)
    fr = synth.mot_frame(M=M, T=T, budget=budget, D=D, seed=M + T)
    samples = fr["samples"]
    gal = ops.Gallery(T + 2, budget, D)
    keep = [t for t in range(T) if t % 7 != 3]  # some tracks never fitted: empty slot -> 1e5 rows
    gal.partial_fit(np.concatenate([samples[t] for t in keep], 0), np.repeat(keep, [len(samples[t]) for t in keep]))
    slots = np.arange(T, dtype=np.int32)
    if T > 2:
        slots[1] = -1  # track without a gallery slot
    args = (gal, fr["det_feat"], fr["det_tlwh"], slots, fr["trk_mean"], fr["trk_SI"].reshape(T, 16), fr["trk_tsu"], 0.995, 0.2, 0.7)
    app, raw, iou = [x.cpu().numpy() for x in ops.assoc_frame(*args, want_raw=True)]
    app_s, raw_s, iou_s = [x.cpu().numpy() for x in ops.assoc_frame(*args, want_raw=True, simt=True)]
    assert np.array_equal(iou, iou_s)
    np.testing.assert_allclose(raw, raw_s, rtol=1e-4, atol=1e-5)  # north_star: distances within 1e-4 relative
    np.testing.assert_allclose(app, app_s, rtol=1e-4, atol=1e-5)
    assert np.array_equal(app > 1e4, app_s > 1e4)
    empty = [t for t in range(T) if t not in keep or slots[t] < 0]
    assert (raw[empty] == 1e5).all()
    live = [t for t in range(T) if t in keep and slots[t] >= 0]
    oraw = O.nn_distance_matrix([samples[t] for t in live], fr["det_feat"])
    np.testing.assert_allclose(raw[live], oraw, rtol=1e-4, atol=1e-5)
    oiou = O.iou_cost([O.track_to_tlwh(m[:4]) for m in fr["trk_mean"]], fr["trk_tsu"], fr["det_tlwh"])
    assert np.array_equal(iou, oiou)


def test_empty_inputs(rb):
    from reidb200 import iou_matching as IM, linear_assignment as LA, matching as MT, tracker as TR

    metric = MT.NearestNeighborDistanceMetric("cosine", 0.2, 5)
    m, ut, ud = TR.match_arrays(metric, np.zeros((0, 8), np.float32), np.zeros((0, 4), np.float32), [], [], np.zeros((0, 8)),
                                np.zeros((0, 16)), [], 0.995, 0.7)
    assert m == [] and ut == [] and ud == []
    assert LA.min_cost_matching(None, 0.2, [], [], [], [])[0] == []
    assert IM.iou(np.array([0., 0, 1, 1]), np.zeros((0, 4))).shape == (0,)


def test_frame_context_equals_frame_costs(rb, monkeypatch):
    """tracker.FrameContext (two C calls per frame from host arrays) returns the crops of preprocess_frame and the cost
    matrices of frame_costs + costs_to_host, bit for bit, over frames of different sizes through the same context."""
    from reidb200 import matching as MT, preprocess, synth, tracker as TR

    monkeypatch.setenv("REIDB200_ASSOC_V2", "1")
    for seed, (M, T) in enumerate([(30, 100), (200, 300), (7, 12)]):
        fr = synth.mot_frame(M=M, T=T, budget=100, D=128, seed=20 + seed)
        metric = MT.NearestNeighborDistanceMetric("cosine", 0.2, 100, max_tracks=T)
        metric.partial_fit(np.concatenate(fr["samples"], 0), np.repeat(np.arange(T), [len(s) for s in fr["samples"]]), list(range(T)))
        fc = TR.FrameContext(metric, max_dets=256, max_tracks=512)
        H, W = fr["img"].shape[:2]
        xyxy = preprocess.xywh_to_xyxy(fr["det_xywh"], W, H)
        xyxy = xyxy[(xyxy[:, 2] > xyxy[:, 0]) & (xyxy[:, 3] > xyxy[:, 1])]
        frame = torch.from_numpy(fr["img"]).cuda()
        conf = np.arange(T) % 5 != 0
        for rep in range(3):  # the staging buffers alternate
            crops = fc.crops(frame, xyxy)
            assert torch.equal(crops, preprocess.preprocess_frame(frame, xyxy))
            det_feat = torch.from_numpy(fr["det_feat"]).cuda()
            app, iou = fc.costs(det_feat, fr["det_tlwh"], np.arange(T), conf, fr["trk_mean"], fr["trk_SI"], fr["trk_tsu"], 0.995, 0.2, 0.7)
            a2, _, i2 = TR.frame_costs(metric, det_feat, fr["det_tlwh"], np.arange(T), conf, fr["trk_mean"], fr["trk_SI"], fr["trk_tsu"],
                                       0.995, 0.2, 0.7)
            assert np.array_equal(app, a2.cpu().numpy()) and np.array_equal(iou, i2.cpu().numpy())
            assert (app[~conf] == 0.2 + 1e-5).all()   # unconfirmed rows: 1e5 before the clamp
        with pytest.raises(ValueError):
            fc.crops(frame, np.array([[5, 5, 5, 40]], np.int32))
    z = fc.costs(det_feat[:0], fr["det_tlwh"][:0], [], [], np.zeros((0, 8)), np.zeros((0, 16)), np.zeros(0, np.int32), 0.995)
    assert z[0].shape == (0, 0)
