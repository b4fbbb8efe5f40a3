"""GPU tests of the reference-named shims through their own call signatures: the per-object Kalman adapter and its
gating_distance, gate_cost_matrix / min_cost_matching / matching_cascade on Track / Detection objects, tracker.match,
NearestNeighborDistanceMetric.distance (+ its unbounded growth and the euclidean restatement), StrongSORT._get_features,
get_feats / get_evaluate_result, index_init_cpu / index_init_gpu with repeated add, Extractor's backbone contract."""
import numpy as np
import pytest
import torch

from oracle import reid_oracle as O

pytestmark = pytest.mark.gpu


def _frames(g):
    for f in range(int(g["n_frames"])):
        yield {k[len("f%d_" % f):]: g[k] for k in g.files if k.startswith("f%d_" % f)}


class _KF:
    def __init__(self, mean, SI):
        self.x, self.SI = np.asarray(mean, np.float64).copy(), np.asarray(SI, np.float64).reshape(4, 4).copy()


class _Track:
    """What the reference's linear_assignment / iou_matching / Tracker._match read of a Track (sort/track.py)."""

    def __init__(self, track_id, mean, SI, tsu, confirmed):
        self.track_id, self.mean, self.time_since_update, self._c = int(track_id), np.asarray(mean, np.float64), int(tsu), bool(confirmed)
        self.kf = _KF(mean, SI)

    def is_confirmed(self):
        return self._c

    def to_tlwh(self):
        ret = self.mean[:4].copy()
        ret[2] *= ret[3]
        ret[:2] -= ret[2:] / 2
        return ret


class _Det:
    def __init__(self, tlwh, feature):
        self.tlwh, self.feature = np.asarray(tlwh, np.float32), np.asarray(feature, np.float32)

    def to_xyah(self):
        ret = self.tlwh.copy()
        ret[:2] += ret[2:] / 2
        ret[2] /= ret[3]
        return ret


def _objects(fr):
    tracks = [_Track(i, m, s, t, c) for i, m, s, t, c in zip(fr["trk_id"], fr["trk_mean"], fr["trk_SI"], fr["trk_tsu"], fr["trk_confirmed"])]
    dets = [_Det(t, f) for t, f in zip(fr["det_tlwh"], fr["det_feat"])]
    return tracks, dets


def _metric_for(fr, budget, max_dist, **kw):
    from reidb200 import matching as MT

    metric = MT.NearestNeighborDistanceMetric("cosine", max_dist, budget, **kw)
    ids = fr["sample_ids"].tolist()
    if ids:
        metric.partial_fit(fr["sample_rows"], np.repeat(fr["sample_ids"], fr["sample_cnt"]), ids)
    return metric


def test_distance_gate_cost_matrix_and_matching_on_reference_objects(rb, golden):
    """The recorded reference Tracker sequence replayed through the object-level shims, the way Tracker._match's
    gated_metric does (sort/tracker.py:176-189): metric.distance -> gate_cost_matrix (in place AND returned) ->
    matching_cascade; then iou_matching.iou_cost -> min_cost_matching; and tracker.match in one call."""
    from reidb200 import iou_matching as IM, linear_assignment as LA, tracker as TR

    g = golden("assoc_seq")
    max_dist, max_iou, lam, budget = (float(g["cfg_max_dist"]), float(g["cfg_max_iou_dist"]), float(g["cfg_mc_lambda"]),
                                      int(g["cfg_budget"]))
    checked = 0
    for fr in _frames(g):
        tracks, dets = _objects(fr)
        metric = _metric_for(fr, budget, max_dist)
        conf = [i for i, t in enumerate(tracks) if t.is_confirmed()]
        if conf and dets:
            feats = np.array([d.feature for d in dets])
            targets = np.array([tracks[i].track_id for i in conf])
            cost = metric.distance(feats, targets)                      # E2 through its own name
            assert cost.dtype == np.float64 and cost.shape == (len(conf), len(dets))
            np.testing.assert_allclose(cost, fr["app_raw"], rtol=1e-4, atol=1e-5)
            ret = LA.gate_cost_matrix(cost, tracks, dets, conf, list(range(len(dets))), lam)   # E7
            assert ret is cost                                           # mutated in place and returned
            np.testing.assert_allclose(cost, fr["app_gated"], rtol=1e-4, atol=1e-5)
            assert np.array_equal(cost > 1e4, fr["app_gated"] > 1e4)
            with pytest.raises(ValueError):                              # the reference's only_position path always raises
                LA.gate_cost_matrix(cost.copy(), tracks, dets, conf, list(range(len(dets))), lam, only_position=True)
            checked += 1

        def gated_metric(trks, ds, track_indices, detection_indices):
            f = np.array([ds[i].feature for i in detection_indices])
            t = np.array([trks[i].track_id for i in track_indices])
            c = metric.distance(f, t)
            return LA.gate_cost_matrix(c, trks, ds, track_indices, detection_indices, lam)

        unconf = [i for i, t in enumerate(tracks) if not t.is_confirmed()]
        m_a, um_a, um_d = LA.matching_cascade(gated_metric, max_dist, 70, tracks, dets, conf)
        cand = unconf + [k for k in um_a if tracks[k].time_since_update == 1]
        um_a = [k for k in um_a if tracks[k].time_since_update != 1]
        m_b, um_b, um_d = LA.min_cost_matching(IM.iou_cost, max_iou, tracks, dets, cand, um_d)
        matches = sorted(m_a + m_b)
        assert np.array_equal(np.array(matches, np.int64).reshape(-1, 2), fr["matches"])
        assert sorted(set(um_a + um_b)) == fr["unmatched_tracks"].tolist() and sorted(um_d) == fr["unmatched_dets"].tolist()
        m, ut, ud = TR.match(metric, tracks, dets, lam, max_iou)         # the object form of the fused path
        assert np.array_equal(np.array(sorted(m), np.int64).reshape(-1, 2), fr["matches"])
        assert sorted(ut) == fr["unmatched_tracks"].tolist() and sorted(ud) == fr["unmatched_dets"].tolist()
    assert checked >= 4


def test_kalman_adapter_objects_replay_reference_sequence(rb, golden):
    """StrongSortKalmanFilterAdapter per-track objects (initiate / predict / update / gating_distance) against the
    recorded reference adapter sequence and the oracle's Mahalanobis distance."""
    from reidb200.strongsort_kf_adapter import StrongSortKalmanFilterAdapter

    g = golden("kalman")
    T, frames = g["z0"].shape[0], g["matched"].shape[0]
    kfs, state = [], []
    for t in range(T):
        kf = StrongSortKalmanFilterAdapter()
        x, P = kf.initiate(g["z0"][t].copy())
        kfs.append(kf)
        state.append((x, P))
        assert np.array_equal(kf.SI, np.zeros((4, 4)))
        assert np.array_equal(kf.gating_distance(np.array([[1., 2., 0.5, 30.]])), np.zeros(1))  # fresh track: SI = 0 (F6)
    rng = np.random.default_rng(3)
    for f in range(frames):
        for t in range(T):
            x, P = kfs[t].predict(*state[t])
            state[t] = (x, P)
            if g["matched"][f, t]:
                x, P = kfs[t].update(state[t][0], state[t][1], g["meas"][f, t], g["conf"][f, t])
                state[t] = (x, P)
        np.testing.assert_allclose(np.stack([s[0] for s in state]), g["x"][f], rtol=1e-8, atol=1e-9)
        np.testing.assert_allclose(np.stack([s[1] for s in state]), g["P"][f], rtol=1e-7, atol=1e-9)
        np.testing.assert_allclose(np.stack([k.SI for k in kfs]), g["SI"][f], rtol=1e-7, atol=1e-12)
        meas = (g["x"][f][:, :4].mean(0) + rng.normal(0, [5, 5, 0.02, 5], (6, 4))).astype(np.float32)
        for t in (0, T // 2, T - 1):
            got = kfs[t].gating_distance(meas)
            want = O.gating_distance(kfs[t].x[:4], kfs[t].SI, meas.astype(np.float64))
            assert got.shape == (6,)
            np.testing.assert_allclose(got, want, rtol=1e-10, atol=1e-12)
    with pytest.raises(ValueError):
        kfs[0].gating_distance(meas, only_position=True)
    with pytest.raises(ValueError):
        kfs[0].gating_distance(meas[:1], only_position=True)
    assert kfs[0].gating_distance(np.zeros((0, 4))).shape == (0,)


def test_metric_is_unbounded_like_the_reference(rb):
    """More live targets than max_tracks and, with budget=None, more samples than the initial ring: the gallery grows and
    every distance equals the one of a gallery that was large enough from the start."""
    from reidb200 import matching as MT

    rng = np.random.default_rng(4)
    D, n_t = 32, 11
    small = MT.NearestNeighborDistanceMetric("cosine", 0.2, None, max_tracks=4, max_unbounded=4)
    big = MT.NearestNeighborDistanceMetric("cosine", 0.2, None, max_tracks=64, max_unbounded=64)
    ref = {}
    for step in range(9):
        n = int(rng.integers(3, 12))
        feats = rng.standard_normal((n, D)).astype(np.float32)
        targets = rng.integers(0, n_t, n)
        active = list(range(n_t)) if step != 5 else [t for t in range(n_t) if t % 3]
        for m in (small, big):
            m.partial_fit(feats, targets, active)
        for f, t in zip(feats, targets):
            ref.setdefault(int(t), []).append(f)
        ref = {k: ref[k] for k in active if k in ref}
    assert small._slots_cap > 4 and small._ring > 4
    q = rng.standard_normal((7, D)).astype(np.float32)
    ids = sorted(ref)
    want = O.nn_distance_matrix([np.stack(ref[k]) for k in ids], q)
    for m in (small, big):
        np.testing.assert_allclose(m.distance(q, ids), want, rtol=1e-4, atol=1e-5)
    assert np.array_equal(small.distance(q, ids), big.distance(q, ids))  # growth moved the rows bit for bit
    for k in ids:
        assert len(small.samples[k]) == len(ref[k])  # budget=None keeps every sample


def test_euclidean_metric_restatement(rb):
    """metric="euclidean" (broken in the reference): smallest squared distance to a stored sample."""
    from reidb200 import matching as MT

    rng = np.random.default_rng(6)
    m = MT.NearestNeighborDistanceMetric("euclidean", 1.0, 3)
    ref = {}
    for step in range(4):
        feats = rng.standard_normal((6, 16)).astype(np.float32)
        targets = rng.integers(0, 3, 6)
        m.partial_fit(feats, targets, [0, 1, 2])
        for f, t in zip(feats, targets):
            ref.setdefault(int(t), []).append(f)
            ref[int(t)] = ref[int(t)][-3:]
    q = rng.standard_normal((5, 16)).astype(np.float32)
    ids = sorted(ref)
    got = m.distance(q, ids)
    want = np.stack([O.pdist(np.stack(ref[k]), q).min(0) for k in ids])
    np.testing.assert_allclose(got, want, rtol=1e-4, atol=1e-5)
    x, y = rng.standard_normal((4, 16)).astype(np.float32), q
    np.testing.assert_allclose(MT._nn_euclidean_distance(x, y), O.pdist(x, y).min(0), rtol=1e-4, atol=1e-5)
    assert sorted(m.samples.keys()) == ids and len(m.samples[ids[0]]) == len(ref[ids[0]])


def test_strongsort_get_features_and_box_helpers(rb, golden):
    """StrongSORT._get_features(bbox_xywh, img): crops by the reference's box rule, one launch, handed to the model --
    against the real PIL / cv2 outputs of the preprocess fixture."""
    from reidb200.strong_sort import StrongSORT

    g = golden("preprocess")
    seen = {}

    def model(batch):
        seen["batch"] = batch
        return batch.flatten(1)[:, :8].cpu().numpy()

    ss = StrongSORT(model, preprocess_mode="cv_linear")
    feats = ss._get_features(g["boxes_xywh"], g["img"])
    assert feats.shape == (len(g["boxes_xywh"]), 8) and ss.width == g["img"].shape[1] and ss.height == g["img"].shape[0]
    batch = seen["batch"].cpu().numpy()
    assert batch.shape == (len(g["boxes_xywh"]), 3, 256, 128)
    np.testing.assert_allclose(batch[g["cv_full_idx"]], g["cv_full"], rtol=0, atol=4.8e-7)   # real cv2 output, 1 ulp
    np.testing.assert_allclose(batch[:, :, ::3, ::3], g["cv_sub"], rtol=0, atol=4.8e-7)
    for b, want in zip(g["boxes_xywh"], g["xyxy"]):
        assert ss._xywh_to_xyxy(b) == tuple(int(v) for v in want)
    ss2 = StrongSORT(model)                                   # boxmot front-end: PIL antialias + ImageNet normalisation
    ss2._get_features(torch.from_numpy(g["boxes_xywh"]), g["img"])
    np.testing.assert_allclose(seen["batch"][0].cpu().numpy(), g["pil0"], rtol=0, atol=1e-6)
    assert ss._get_features(np.zeros((0, 4), np.float32), g["img"]).shape == (0,)
    tlwh = StrongSORT._xywh_to_tlwh(g["boxes_xywh"])
    np.testing.assert_array_equal(tlwh[:, 0], g["boxes_xywh"][:, 0] - g["boxes_xywh"][:, 2] / 2.0)
    assert ss._tlwh_to_xyxy((5.7, -3.2, 1000.0, 20.9)) == (5, 0, ss.width - 1, 17) and ss._xyxy_to_tlwh((1, 2, 11, 22)) == (1, 2, 10, 20)


def test_preprocess_clips_like_numpy_and_extractor_needs_a_backbone(rb):
    from reidb200 import feature_extractor as FE, preprocess as P

    rng = np.random.default_rng(2)
    img = rng.integers(0, 256, (60, 80, 3), dtype=np.uint8)
    xyxy = np.array([[10, 5, 200, 100], [70, 50, 80, 60]])       # beyond the frame: img[y1:y2, x1:x2] clips silently
    got = P.preprocess_frame(img, xyxy).cpu().numpy()
    want = O.preprocess_deepsort([img[5:100, 10:200], img[50:60, 70:80]])
    assert np.array_equal(got, want)
    with pytest.raises(ValueError):
        P.preprocess_frame(img, np.array([[-3, 0, 10, 10]]))
    with pytest.raises(ValueError):
        P.preprocess_frame(img, np.array([[90, 0, 95, 10]]))     # entirely outside: empty crop
    with pytest.raises(ValueError):
        FE.Extractor("weights.t7")                               # no backbone: never silently random weights
    with pytest.raises(ValueError):
        FE.Extractor()
    ex = FE.Extractor(standin=True)
    out = ex([img[5:50, 10:40], img[0:30, 0:20]])
    assert out.shape == (2, 512)
    net = torch.nn.Sequential(torch.nn.AdaptiveAvgPool2d(1), torch.nn.Flatten())
    assert FE.Extractor(net=net)([img[5:50, 10:40]]).shape == (1, 3)


def test_get_feats_and_get_evaluate_result(rb):
    """reid/evaluate.py:5-27 on a stand-in model and loaders: the concatenated features feed evaluate_all."""
    from reidb200 import evaluate as E, synth

    d = synth.market_like(Q=40, G=300, D=32, n_ids=60, seed=9, noise=1.2)

    class Model(torch.nn.Module):
        def forward(self, image, cam):
            return None, image.flatten(1)   # (logits, embedding): get_feats keeps element [1]

    def loader(f, l, c, bs=16):
        return [(torch.from_numpy(f[i:i + bs]).reshape(-1, 1, 4, 8), torch.from_numpy(l[i:i + bs]), torch.from_numpy(c[i:i + bs]))
                for i in range(0, len(l), bs)]

    model = Model()
    labels, feats, cams = E.get_feats(model, loader(d["gf"], d["gl"], d["gc"]))
    assert torch.equal(labels, torch.from_numpy(d["gl"])) and torch.equal(cams, torch.from_numpy(d["gc"]))
    assert feats.is_cuda and torch.equal(feats.cpu(), torch.from_numpy(d["gf"]))
    assert model.training
    cmc, ap = E.get_evaluate_result(model, loader(d["qf"], d["ql"], d["qc"]), loader(d["gf"], d["gl"], d["gc"]))
    ocmc, oap = O.evaluate_all(d["qf"], d["ql"], d["qc"], d["gf"], d["gl"], d["gc"])
    assert np.abs(cmc.numpy() - ocmc).max() <= 1e-6 and abs(ap - oap) <= 1e-6 and ocmc[0] < 0.95


def test_index_init_cpu_and_repeated_add(rb):
    """index_init_cpu (IndexFlatL2 stand-in) and index_init_gpu (IndexShards) with add() called more than once: ids keep
    counting in call order; same neighbours as one add of the concatenation and as the oracle."""
    from reidb200 import faiss_utils as F

    rng = np.random.default_rng(8)
    xb = rng.standard_normal((900, 48)).astype(np.float32)
    xq = rng.standard_normal((25, 48)).astype(np.float32)
    Do, Io = O.knn_l2(xb, xq, 7)
    for make in (lambda: F.index_init_cpu(48), lambda: F.index_init_gpu(1, 48), lambda: F.index_init_gpu(0, 48)):
        idx = make()
        idx.add(xb[:400])
        idx.add(xb[400:650])
        idx.add(xb[650:])
        assert idx.ntotal == 900
        D, I = idx.search(xq, 7)
        assert I.dtype == np.int64 and np.array_equal(I, Io)
        np.testing.assert_allclose(D, Do, rtol=1e-4, atol=1e-5)
        Dt, It = F.search_index_pytorch(idx, torch.from_numpy(xq).cuda(), 7)
        assert np.array_equal(It.cpu().numpy(), Io)
        idx.reset()
        assert idx.ntotal == 0
        idx.add(xb[:50])
        assert np.array_equal(idx.search(xq, 3)[1], O.knn_l2(xb[:50], xq, 3)[1])


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_index_shards_over_two_gpus_in_one_process(rb):
    """index_init_gpu(2, d): the one-process IndexShards drop-in touches a second device, where every >48 KB
    dynamic-shared-memory kernel needs its own opt-in (per-device function attributes)."""
    from reidb200 import faiss_utils as F, ops

    rng = np.random.default_rng(9)
    xb = rng.standard_normal((5000, 64)).astype(np.float32)
    xq = rng.standard_normal((300, 64)).astype(np.float32)
    idx = F.index_init_gpu(2, 64)
    assert idx.ngpus == 2
    idx.add(xb[:3000])
    idx.add(xb[3000:])
    D, I = idx.search(xq, 20)
    D1, I1 = ops.knn(xq, xb, 20, "sql2")
    assert np.array_equal(I, I1.cpu().numpy()) and np.array_equal(D, D1.cpu().numpy())
    with torch.cuda.device(1):  # the re-ranking kernels on the second device too
        J = ops.jaccard(torch.from_numpy(xb[:1500]).cuda(1), 20, 6)
    J0 = ops.jaccard(torch.from_numpy(xb[:1500]).cuda(0), 20, 6)
    assert torch.equal(J.cpu(), J0.cpu())


def test_dbscan_refuses_an_asymmetric_matrix(rb):
    from reidb200 import cluster

    rng = np.random.default_rng(1)
    d = rng.random((50, 50)).astype(np.float32)
    with pytest.raises(ValueError):
        cluster.DBSCAN(eps=0.3, min_samples=3).fit_predict(d)
    with pytest.raises(ValueError):
        cluster.DBSCAN(eps=0.3, min_samples=3).fit_predict(torch.from_numpy(d).cuda())
    s = (d + d.T) / 2
    np.fill_diagonal(s, 0)
    lab = cluster.DBSCAN(eps=0.3, min_samples=3).fit_predict(s)
    olab, _ = O.dbscan_precomputed(s, 0.3, 3)
    assert np.array_equal(lab, olab)
