"""Pin the oracle against the reference's own Python code and write tests/golden/*.npz.

TEST INFRASTRUCTURE ONLY.  Runs only in the build container (needs /root/reference):

    python -m oracle.make_golden

For every stage of the hot path it (1) runs the *reference* function, imported in place by
oracle/ref_import.py, on seeded inputs, (2) runs the numpy restatement in oracle/reid_oracle.py
on the same inputs, (3) asserts agreement (bit-exact for integer/index work, tight fp
tolerances otherwise) and (4) stores inputs + reference outputs as a small fixture.  The
fixtures travel to the GPU box; /root/reference does not.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_import, reid_oracle as O  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def _save(name, **arrs):
    os.makedirs(GOLD, exist_ok=True)
    path = os.path.join(GOLD, name + ".npz")
    np.savez_compressed(path, **arrs)
    print("  wrote %-28s %7.1f KB" % (name + ".npz", os.path.getsize(path) / 1024))


def _unit(x):
    return (x / np.linalg.norm(x, axis=1, keepdims=True)).astype(np.float32)


# ---------------------------------------------------------------- B1/B2 distances
def gold_dist():
    U = ref_import.load_file("ref_losses_utils", "reid/losses/utils.py")
    rng = np.random.default_rng(0)
    x = rng.standard_normal((37, 48)).astype(np.float32)
    y = rng.standard_normal((53, 48)).astype(np.float32)
    y[5] = x[3]  # a duplicate row: exercises the clamp(1e-12) branch
    e = U.euclidean_dist(torch.from_numpy(x), torch.from_numpy(y)).numpy()
    c = U.cosine_dist(torch.from_numpy(x), torch.from_numpy(y)).numpy()
    np.testing.assert_allclose(O.euclidean_dist(x, y), e, rtol=2e-5, atol=1e-2)  # the duplicate pair: sqrt of fp32 cancellation noise
    mask = e > 1e-2
    np.testing.assert_allclose(O.euclidean_dist(x, y)[mask], e[mask], rtol=1e-5)
    np.testing.assert_allclose(O.cosine_dist(x, y), c, rtol=0, atol=2e-7)
    _save("dist", x=x, y=y, euclid=e, cosine=c)


# ---------------------------------------------------------------- B3-B6 evaluation
def ranked_eval(E, dist, ql, qc, gl, gc):
    """Reference compute_mAP (reid/evaluate.py:78-105) over every row of a DISTANCE matrix (ascending, stable argsort)
    with the reference's good / junk rule (:66-72).  Returns (CMC f32[G] / Q, mAP, per-query AP, first-hit rank (-1 = no
    valid match), margin) -- margin[i] = smallest relative gap between a good item and a non-good neighbour in query i's
    junk-free ranking: where it is below the distance tolerance, a swap of the two is a tie, not an error."""
    nq, ng = dist.shape
    CMC = torch.IntTensor(ng).zero_()
    ap_sum = 0.0
    aps = np.zeros(nq)
    first = np.full(nq, -1, np.int32)
    margin = np.full(nq, np.inf)
    for i in range(nq):
        index = np.argsort(dist[i], kind="stable")
        query_index = np.argwhere(gl == ql[i])
        camera_index = np.argwhere(gc == qc[i])
        good_index = np.setdiff1d(query_index, camera_index, assume_unique=True)
        junk_index1 = np.argwhere(gl == -1)
        junk_index2 = np.intersect1d(query_index, camera_index)
        junk_index = np.append(junk_index2, junk_index1)
        ap, cmc = E.compute_mAP(index, good_index, junk_index)
        if cmc[0] == -1:
            continue
        CMC = CMC + cmc
        ap_sum += ap
        aps[i] = ap
        first[i] = int(np.argmax(cmc.numpy()))
        keep = index[np.isin(index, junk_index, invert=True)]
        dv = dist[i][keep].astype(np.float64)
        isg = np.isin(keep, good_index)
        edge = isg[1:] != isg[:-1]
        if edge.any():
            margin[i] = ((dv[1:] - dv[:-1])[edge] / np.maximum(np.abs(dv[1:][edge]), 1e-12)).min()
    return (CMC.float() / nq).numpy(), ap_sum / nq, aps, first, margin


def gold_eval():
    """HARD synthetic data (synth.hard_noise): first hits spread over many ranks, CMC is a curve, APs take many values --
    a kernel that mis-ranks anything but the top item fails against this fixture."""
    E = ref_import.load_file("ref_evaluate", "reid/evaluate.py")
    sys.path.insert(0, os.path.join(ROOT, "real-time-reid-tracking_b200"))
    import synth  # noqa

    d = synth.market_like(Q=90, G=600, D=32, n_ids=150, seed=3, orphan_frac=0.05, noise=1.2)
    t = {k: torch.from_numpy(v) for k, v in d.items()}
    import contextlib, io
    with contextlib.redirect_stdout(io.StringIO()):
        cmc, ap = E.evaluate_all(t["qf"], t["ql"], t["qc"], t["gf"], t["gl"], t["gc"])
    cmc = cmc.numpy()
    ocmc, oap = O.evaluate_all(d["qf"], d["ql"], d["qc"], d["gf"], d["gl"], d["gc"])
    assert np.array_equal(ocmc, cmc), "CMC mismatch"
    assert abs(oap - ap) < 1e-12, (oap, ap)
    # per-query AP and first-hit rank (reference evaluate())
    aps, first = [], []
    for i in range(len(d["ql"])):
        a, c = E.evaluate(t["qf"][i], t["ql"][i], t["qc"][i], t["gf"], t["gl"], t["gc"])
        c = c.numpy()
        aps.append(float(a))
        first.append(-1 if c[0] == -1 else int(np.argmax(c)))
    first = np.array(first, np.int32)
    assert first.max() >= 5 and len(np.unique(first)) >= 5 and (first < 0).any(), "fixture too easy: %s" % np.bincount(first[first >= 0])
    assert len(np.unique(np.round(aps, 9))) > 40 and cmc[0] < 0.9 < cmc[-1]
    print("  eval: rank-1 %.4f rank-5 %.4f mAP %.4f, first hits up to rank %d" % (cmc[0], cmc[4], ap, first.max()))
    _save("eval", **d, cmc=cmc, ap=np.float64(ap), ap_per_query=np.array(aps), first_hit=first)


def gold_rerank_eval():
    """Re-ranked evaluation at a size the GPU tests compare entry by entry: the reference's compute_jaccard_distance on
    merged = cat(gallery, query) (image_reid_inference.py:270), blended with the reference's euclidean_dist (the blend
    itself is the north_star addition), ranked and reduced by the reference's compute_mAP."""
    E = ref_import.load_file("ref_evaluate", "reid/evaluate.py")
    F = ref_import.load_file("ref_faiss_utils", "reid/faiss_utils.py")
    U = ref_import.load_file("ref_losses_utils", "reid/losses/utils.py")
    sys.path.insert(0, os.path.join(ROOT, "real-time-reid-tracking_b200"))
    import synth  # noqa

    Q, G, lam = 160, 900, 0.3
    d = synth.market_like(Q=Q, G=G, D=64, n_ids=150, seed=8, orphan_frac=0.04, noise=1.4)
    t = {k: torch.from_numpy(v) for k, v in d.items()}
    merged = torch.cat([t["gf"], t["qf"]], 0)
    out = dict(d)
    for tag, k1, k2 in (("k20_6", 20, 6), ("k12_3", 12, 3)):
        J = F.compute_jaccard_distance(merged, k1=k1, k2=k2, print_flag=False, search_option=3)
        Jo = O.compute_jaccard_distance(merged.numpy(), k1, k2)
        np.testing.assert_allclose(Jo, J, rtol=0, atol=2e-6)
        orig = U.euclidean_dist(t["qf"], t["gf"]).numpy()
        dist = (J[G:, :G] * np.float32(1.0 - lam) + orig * np.float32(lam)).astype(np.float32)
        np.testing.assert_allclose(O.re_ranking_from_feats(d["qf"], d["gf"], k1, k2, lam), dist, rtol=0, atol=2e-6)
        cmc, mAP, aps, first, margin = ranked_eval(E, dist, d["ql"], d["qc"], d["gl"], d["gc"])
        ocmc, oap = O.evaluate_all_from_dist(dist, d["ql"], d["qc"], d["gl"], d["gc"])
        assert np.array_equal(ocmc, cmc) and abs(oap - mAP) < 1e-12
        assert first.max() >= 3 and cmc[0] < 0.995, "re-ranked fixture too easy"
        print("  rerank_eval %s: rank-1 %.4f mAP %.4f, first hits up to rank %d, %d queries with margin < 1e-5"
              % (tag, cmc[0], mAP, first.max(), (margin < 1e-5).sum()))
        out.update({"dist_" + tag: dist, "cmc_" + tag: cmc, "map_" + tag: np.float64(mAP), "ap_q_" + tag: aps,
                    "first_" + tag: first, "margin_" + tag: margin})
    _save("rerank_eval", lambda_value=np.float64(lam), **out)


# ---------------------------------------------------------------- D1-D7 Jaccard re-ranking
def gold_jaccard():
    F = ref_import.load_file("ref_faiss_utils", "reid/faiss_utils.py")
    rng = np.random.default_rng(5)
    N, D = 260, 32
    cent = rng.standard_normal((30, D)).astype(np.float32)
    feats = _unit(cent[rng.integers(0, 30, N)] + 0.5 * rng.standard_normal((N, D)).astype(np.float32))
    out = {}
    for tag, k1, k2, f16 in (("k20_6", 20, 6, False), ("k10_1", 10, 1, False), ("k20_6_f16", 20, 6, True)):
        J = F.compute_jaccard_distance(torch.from_numpy(feats), k1=k1, k2=k2, print_flag=False,
                                       search_option=3, use_float16=f16)
        Jo = O.compute_jaccard_distance(feats, k1, k2, f16)
        assert J.dtype == Jo.dtype and J.shape == Jo.shape
        np.testing.assert_allclose(Jo.astype(np.float32), J.astype(np.float32), rtol=0, atol=2e-3 if f16 else 2e-6)
        out["J_" + tag] = J
    _, rank = O.knn_l2(feats, feats, 20)
    # k_reciprocal_neigh on a few rows
    for i in (0, 17, 255):
        assert np.array_equal(F.k_reciprocal_neigh(rank, i, 20), O.k_reciprocal_neigh(rank, i, 20))
        assert np.array_equal(F.k_reciprocal_neigh(rank, i, 10), O.k_reciprocal_neigh(rank, i, 10))
    # sparse rows restatement agrees with the dense reference V path (through J): covered above.
    _save("jaccard", feats=feats, initial_rank=rank.astype(np.int32), **out)


# ---------------------------------------------------------------- A1-A3 crop / resize / normalise
def gold_preprocess():
    FE = None
    # feature_extractor.py imports a backbone that does not exist: stub the module tree
    for name in ("reid", "reid.backbones", "reid.backbones.SERes18_IBN"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["reid.backbones.SERes18_IBN"].SEDense18_IBN = object
    FE = ref_import.load_file("ref_feature_extractor", "modification_deepsort/feature_extractor.py")
    for name in ("reid.backbones.SERes18_IBN", "reid.backbones", "reid"):
        sys.modules.pop(name, None)
    ex = object.__new__(FE.Extractor)
    import torchvision.transforms as transforms
    ex.size = (128, 256)
    ex.norm = transforms.Compose([transforms.ToTensor(), transforms.Normalize([0.5, 0.5, 0.5], [0.5, 0.5, 0.5])])

    rng = np.random.default_rng(7)
    H, W = 270, 480
    img = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    # smooth part so that interpolation is exercised on structured data too
    yy, xx = np.mgrid[0:H, 0:W]
    img[:, : W // 2, 0] = ((np.sin(xx[:, : W // 2] / 9.0) * 0.5 + 0.5) * 255).astype(np.uint8)
    # centre-xywh boxes: upscale, downscale, clamped at borders, 1-pixel-wide, identity size
    boxes_xywh = np.array([[100.3, 120.7, 40.2, 90.9], [240.0, 135.0, 300.0, 260.0], [5.0, 8.0, 30.0, 40.0],
                           [470.0, 260.0, 40.0, 60.0], [200.5, 100.5, 128.0, 256.0], [300.0, 50.0, 1.2, 37.0],
                           [64.0, 128.0, 128.5, 256.5]], np.float32)

    class _S:  # StrongSORT._xywh_to_xyxy needs width/height attributes only
        pass
    SS = ref_import.load_file_pkg("boxmot.trackers.strongsort.strong_sort")
    s = _S()
    s.width, s.height = W, H
    xyxy = np.array([SS.StrongSORT._xywh_to_xyxy(s, b) for b in boxes_xywh], np.int32)
    oxyxy = np.array([O.xywh_to_xyxy(b, W, H) for b in boxes_xywh], np.int32)
    assert np.array_equal(xyxy, oxyxy)
    crops = [img[y1:y2, x1:x2] for x1, y1, x2, y2 in xyxy]
    ref_cv = ex._preprocess(crops).numpy()
    ocv = O.preprocess_deepsort(crops)
    err = np.abs(ocv - ref_cv).max()
    assert err <= 4.8e-7, err  # 1 ulp of the [0,1] image scaled by 1/std=2
    # boxmot PIL path
    RM = ref_import.load_file_pkg("boxmot.appearance.reid_multibackend")
    import torchvision.transforms as T
    rm = object.__new__(RM.ReIDDetectMultiBackend)
    torch.nn.Module.__init__(rm)
    rm.device = torch.device("cpu")
    rm.preprocess = T.Compose([T.Resize((256, 128)), T.ToTensor(),
                               T.Normalize(mean=[0.485, 0.456, 0.406], std=[0.229, 0.224, 0.225])])
    rm.to_pil = T.ToPILImage()
    ref_pil = rm._preprocess(crops).numpy()
    opil = O.preprocess_boxmot(crops)
    err2 = np.abs(opil - ref_pil).max()
    assert err2 <= 1e-6, err2
    # the uint8 resample must be bit-exact
    from PIL import Image
    for c in crops:
        want = np.asarray(Image.fromarray(c).resize((128, 256), Image.BILINEAR))
        assert np.array_equal(O.resize_pil_bilinear_u8(c, 256, 128), want)
    pil_u8 = np.stack([np.asarray(Image.fromarray(c).resize((128, 256), Image.BILINEAR)) for c in crops])
    # the PIL path is stored as its bit-exact uint8 resample plus the normalised float of crop 0
    _save("preprocess", img=img, boxes_xywh=boxes_xywh, xyxy=xyxy,
          cv_full_idx=np.array([0, 1, 5], np.int32), cv_full=ref_cv[[0, 1, 5]].astype(np.float32),
          cv_sub=ref_cv[:, :, ::3, ::3].astype(np.float32), pil_u8=pil_u8, pil0=ref_pil[0].astype(np.float32))


# ---------------------------------------------------------------- E  association costs
def _mk_tracker(seed=11, frames=9, M=9, D=24, max_dist=0.2, max_iou_dist=0.7):
    """Run the reference StrongSORT Tracker on a synthetic sequence and record, per frame, the
    state that feeds _match() and what _match() returned."""
    TR = ref_import.load_file_pkg("boxmot.trackers.strongsort.sort.tracker")
    DET = ref_import.load_file_pkg("boxmot.trackers.strongsort.sort.detection")
    MT = ref_import.load_file_pkg("boxmot.utils.matching")
    rng = np.random.default_rng(seed)
    metric = MT.NearestNeighborDistanceMetric("cosine", max_dist, 5)
    trk = TR.Tracker(metric, max_iou_dist=max_iou_dist, max_age=4, n_init=2, mc_lambda=0.995, ema_alpha=0.9)
    ids = rng.standard_normal((M, D)).astype(np.float32)
    pos = np.stack([rng.uniform(100, 1700, M), rng.uniform(100, 900, M)], 1)
    hh = rng.uniform(80, 300, M)
    frames_out = []
    for f in range(frames):
        pos += rng.normal(0, 4, pos.shape)
        present = rng.random(M) > 0.2
        if f == 5:
            pos[0] += 400  # a jump: Mahalanobis gate must reject it
        dets = []
        for m in np.where(present)[0]:
            h = hh[m] * rng.uniform(0.97, 1.03)
            w = 0.4 * h
            tlwh = np.array([pos[m, 0] - w / 2, pos[m, 1] - h / 2, w, h], np.float32)
            feat = ids[m] + 0.25 * rng.standard_normal(D).astype(np.float32)
            dets.append(DET.Detection(tlwh, 0.5 + 0.4 * rng.random(), torch.from_numpy(feat.copy())))
        # two clutter detections
        for _ in range(2):
            h = rng.uniform(80, 300)
            tlwh = np.array([rng.uniform(0, 1800), rng.uniform(0, 900), 0.4 * h, h], np.float32)
            dets.append(DET.Detection(tlwh, 0.6, torch.from_numpy(rng.standard_normal(D).astype(np.float32))))
        trk.predict()
        rec = dict(
            det_tlwh=np.stack([d.tlwh for d in dets]).astype(np.float32),
            det_feat=np.stack([d.feature for d in dets]).astype(np.float32),
            trk_id=np.array([t.track_id for t in trk.tracks], np.int64),
            trk_confirmed=np.array([t.is_confirmed() for t in trk.tracks], np.bool_),
            trk_tsu=np.array([t.time_since_update for t in trk.tracks], np.int32),
            trk_mean=np.array([t.mean for t in trk.tracks], np.float64).reshape(-1, 8),
            trk_SI=np.array([t.kf.SI for t in trk.tracks], np.float64).reshape(-1, 4, 4),
        )
        samples = {int(k): np.stack(v).astype(np.float32) for k, v in metric.samples.items()}
        m_, ut, ud = trk._match(dets)
        rec["matches"] = np.array(sorted(m_), np.int64).reshape(-1, 2)
        rec["unmatched_tracks"] = np.array(sorted(ut), np.int64)
        rec["unmatched_dets"] = np.array(sorted(ud), np.int64)
        # cost matrices of both stages on the *full* track/detection sets (what the fused kernel emits)
        conf_idx = [i for i, t in enumerate(trk.tracks) if t.is_confirmed()]
        if conf_idx and dets:
            LA = sys.modules["boxmot.trackers.strongsort.sort.linear_assignment"]
            feats = np.array([d.feature for d in dets])
            cost = metric.distance(feats, np.array([trk.tracks[i].track_id for i in conf_idx]))
            raw = cost.copy()
            cost = LA.gate_cost_matrix(cost, trk.tracks, dets, conf_idx, list(range(len(dets))), trk.mc_lambda)
            rec["app_raw"] = raw
            rec["app_gated"] = cost.copy()
            # oracle check
            oraw = O.nn_distance_matrix([samples[int(trk.tracks[i].track_id)] for i in conf_idx], feats)
            np.testing.assert_allclose(oraw, raw, rtol=0, atol=3e-7)
            meas = O.tlwh_to_xyah_f32(rec["det_tlwh"])
            og = O.gate_cost_matrix(raw.copy(), rec["trk_mean"][conf_idx, :4], rec["trk_SI"][conf_idx], meas, trk.mc_lambda)
            np.testing.assert_allclose(og, cost, rtol=1e-13, atol=0)
        else:
            rec["app_raw"] = np.zeros((0, len(dets)))
            rec["app_gated"] = np.zeros((0, len(dets)))
        IM = sys.modules["boxmot.trackers.strongsort.sort.iou_matching"]
        if trk.tracks and dets:
            ic = IM.iou_cost(trk.tracks, dets)
            oic = O.iou_cost([O.track_to_tlwh(m[:4]) for m in rec["trk_mean"]], rec["trk_tsu"], rec["det_tlwh"])
            np.testing.assert_allclose(oic, ic, rtol=1e-14, atol=0)
            rec["iou_cost"] = ic
        else:
            rec["iou_cost"] = np.zeros((len(trk.tracks), len(dets)))
        rec["sample_ids"] = np.array(sorted(samples), np.int64)
        rec["sample_cnt"] = np.array([len(samples[k]) for k in sorted(samples)], np.int32)
        rec["sample_rows"] = (np.concatenate([samples[k] for k in sorted(samples)], 0)
                              if samples else np.zeros((0, D), np.float32))
        frames_out.append(rec)
        cls = np.zeros(len(dets))
        confs = np.array([d.confidence for d in dets])
        trk.update(dets, cls, confs)
    return frames_out, dict(max_dist=max_dist, max_iou_dist=max_iou_dist, mc_lambda=0.995, budget=5)


def gold_assoc():
    frames, cfg = _mk_tracker()
    flat = {"n_frames": np.int32(len(frames))}
    for k, v in cfg.items():
        flat["cfg_" + k] = np.float64(v)
    n_match = 0
    for f, rec in enumerate(frames):
        for k, v in rec.items():
            flat["f%d_%s" % (f, k)] = v
        n_match += len(rec["matches"])
    assert n_match > 20, "sequence too easy/hard: %d matches" % n_match
    assert any((rec["app_gated"] > 1e4).any() for rec in frames), "gate never fired"
    _save("assoc_seq", **flat)

    # DIoU (modification_deepsort/iou_matching.py) incl. its __main__ example
    DI = ref_import.load_file("ref_diou", "modification_deepsort/iou_matching.py")
    bbox = np.array([10, 12, 8, 9], np.float64)
    rng = np.random.default_rng(2)
    cands = np.concatenate([np.array([[10, 12, 8, 9], [12, 14, 8, 9], [30, 40, 5, 5]], np.float64),
                            rng.uniform(1, 40, (20, 4))])
    want = DI.iou(bbox, cands)
    np.testing.assert_allclose(O.diou(bbox, cands), want, rtol=1e-14)
    IM = ref_import.load_file_pkg("boxmot.trackers.strongsort.sort.iou_matching")
    want_iou = IM.iou(bbox, cands)
    np.testing.assert_allclose(O.iou(bbox, cands), want_iou, rtol=1e-14)
    # matching.py helpers
    MT = ref_import.load_file_pkg("boxmot.utils.matching")
    a = rng.standard_normal((7, 16)).astype(np.float32)
    b = rng.standard_normal((5, 16)).astype(np.float32)
    cd = MT._cosine_distance(a, b)
    pd = MT._pdist(a, b)
    nn = MT._nn_cosine_distance(a, b)
    np.testing.assert_allclose(O.cosine_distance(a, b), cd, atol=2e-7)
    np.testing.assert_allclose(O.pdist(a, b), pd, atol=1e-5)
    np.testing.assert_allclose(O.nn_cosine_distance(a, b), np.asarray(nn), atol=2e-7)
    _save("assoc_misc", bbox=bbox, cands=cands, diou=want, iou=want_iou, a=a, b=b,
          cosine=np.asarray(cd), pdist=np.asarray(pd), nn_cosine=np.asarray(nn))


# ---------------------------------------------------------------- N2 box-pair costs of the other boxmot trackers
def gold_boxpairs():
    IO = ref_import.load_file_pkg("boxmot.utils.iou")
    MT = ref_import.load_file_pkg("boxmot.utils.matching")
    rng = np.random.default_rng(3)

    def boxes(n):
        xy = rng.uniform(0, 1500, (n, 2))
        wh = rng.uniform(5, 300, (n, 2))
        return np.concatenate([xy, xy + wh], 1)

    a64, b64 = boxes(23), boxes(31)
    b64[3] = a64[2]            # identical pair: IoU 1, CIoU's alpha = 0/0 (NaN in the reference too)
    b64[4, :2] += 5000.0       # far away: no overlap
    b64[4, 2:] += 5000.0
    out = {"a64": a64, "b64": b64}
    for dt, tag in ((np.float64, "64"), (np.float32, "32")):
        a, b = a64.astype(dt), b64.astype(dt)
        for mode, fn in (("iou", IO.iou_batch), ("giou", IO.giou_batch), ("diou", IO.diou_batch), ("ciou", IO.ciou_batch)):
            with np.errstate(invalid="ignore"):
                want = fn(a, b)
                got = O.box_pairs(a, b, mode)
            assert want.dtype == dt and np.array_equal(want, got, equal_nan=True), (mode, tag)
            out["%s_%s" % (mode, tag)] = want
    a32, b32 = a64.astype(np.float32), b64.astype(np.float32)
    want = MT.bbox_ious(a32, b32)
    assert np.array_equal(want, O.bbox_ious(a32, b32))
    out["plus_one"] = want
    want = MT.iou_distance(list(a32), list(b32))
    assert want.dtype == np.float32 and np.array_equal(want, O.iou_distance(a32, b32))
    out["iou_distance"] = want

    class Obj:
        def __init__(self, tlbr, score=1.0, feat=None):
            self.tlbr, self.score, self.curr_feat, self.smooth_feat = tlbr, score, feat, feat

    scores = rng.uniform(0.1, 1.0, 31)
    ft, fd = rng.standard_normal((23, 32)).astype(np.float32), rng.standard_normal((31, 32)).astype(np.float32)
    trk = [Obj(a32[i], 1.0, ft[i]) for i in range(23)]
    det = [Obj(b32[i], float(scores[i]), fd[i]) for i in range(31)]
    want = MT.fuse_score(MT.iou_distance(trk, det), det)
    assert want.dtype == np.float64 and np.array_equal(want, O.fuse_score(out["iou_distance"], scores))
    out["fuse_score"] = want
    emb = MT.embedding_distance(trk, det)
    np.testing.assert_allclose(O.embedding_distance(ft, fd), emb, rtol=0, atol=1e-12)
    out["embedding"] = emb
    want = MT.fuse_iou(emb, trk, det)
    assert want.dtype == np.float64 and np.array_equal(want, O.fuse_iou(emb, a32, b32))
    out["fuse_iou64"] = want
    emb32 = emb.astype(np.float32)
    want = MT.fuse_iou(emb32, trk, det)
    assert want.dtype == np.float32 and np.array_equal(want, O.fuse_iou(emb32, a32, b32))
    out["fuse_iou32"] = want
    _save("boxpairs", scores=scores, feat_t=ft, feat_d=fd, **out)


def gold_ocsort():
    """OC-SORT associate: the reference's own iou_matrix / final_cost, captured at its linear_assignment call."""
    A = ref_import.load_file_pkg("boxmot.utils.association")
    from scipy.optimize import linear_sum_assignment

    rng = np.random.default_rng(5)
    T, D = 17, 21
    xy, wh = rng.uniform(0, 900, (T, 2)), rng.uniform(20, 200, (T, 2))
    trk = np.concatenate([xy, xy + wh, np.zeros((T, 1))], 1)
    perm = rng.permutation(T)
    det = np.concatenate([trk[perm, :4] + rng.normal(0, 6, (T, 4)), rng.uniform(0.3, 1, (T, 1))], 1)
    xy, wh = rng.uniform(0, 900, (D - T, 2)), rng.uniform(20, 200, (D - T, 2))
    det = np.concatenate([det, np.concatenate([xy, xy + wh, rng.uniform(0.3, 1, (D - T, 1))], 1)], 0)
    det[1, :4] = trk[perm[0], :4] + 3.0   # two detections on one track: the Hungarian branch is taken
    vel = rng.normal(0, 1, (T, 2))
    vel /= np.linalg.norm(vel, axis=1, keepdims=True)
    vel[3] = 0.0
    prev = np.concatenate([trk[:, :4] - rng.normal(0, 10, (T, 4)), np.ones((T, 1))], 1)
    prev[[2, 9]] = -1.0                   # no previous observation
    captured = {}

    def capture(cost):
        captured["cost"] = np.array(cost)
        x, y = linear_sum_assignment(cost)
        return np.array(list(zip(x, y)))

    A.linear_assignment = capture
    m, ud, ut = A.associate(det, trk, 0.3, vel, prev, 0.2)
    assert "cost" in captured, "fixture must take the assignment branch"
    iou, ang = O.ocsort_cost(det, trk, vel, prev, 0.2)
    Y, X = A.speed_direction_batch(det, prev)
    np.testing.assert_allclose(-(iou + ang), captured["cost"], rtol=0, atol=1e-15)
    _save("ocsort", det=det, trk=trk, vel=vel, prev=prev, vdc=np.float64(0.2), final_cost=captured["cost"], matches=m,
          unmatched_dets=np.asarray(ud), unmatched_trks=np.asarray(ut), Y=Y, X=X)


def gold_kalman():
    """N3: StrongSortKalmanFilterAdapter driven like Track does (one object per track): initiate, then per frame
    predict and, for matched tracks, update with a float32 xyah measurement and a confidence."""
    KA = ref_import.load_file_pkg("boxmot.motion.adapters.strongsort_kf_adapter")
    rng = np.random.default_rng(17)
    T, frames = 9, 7
    z0 = np.stack([rng.uniform(100, 1500, T), rng.uniform(100, 900, T), rng.uniform(0.3, 0.6, T), rng.uniform(60, 300, T)], 1)
    kfs, state = [], []
    for t in range(T):
        kf = KA.StrongSortKalmanFilterAdapter()
        x, P = kf.initiate(z0[t].copy())
        kfs.append(kf)
        state.append([x.copy(), P.copy(), kf.SI.copy()])
        ox, oP, oSI = O.kf_initiate(z0[t])
        assert np.array_equal(ox, x) and np.array_equal(oP, P) and np.array_equal(oSI, kf.SI)
    matched = rng.uniform(size=(frames, T)) < 0.7
    matched[0, 0] = False
    meas = np.zeros((frames, T, 4), np.float32)
    conf = rng.uniform(0.3, 0.95, (frames, T))
    xs, Ps, SIs = [], [], []
    for f in range(frames):
        for t in range(T):
            x, P = kfs[t].predict(state[t][0], state[t][1])
            ox, oP = O.kf_predict(state[t][0], state[t][1])
            np.testing.assert_allclose(ox, x, rtol=1e-13, atol=0)
            np.testing.assert_allclose(oP, P, rtol=1e-12, atol=1e-12)
            state[t][0], state[t][1] = x.copy(), P.copy()
            if matched[f, t]:
                z = (x[:4] + rng.normal(0, [4, 4, 0.01, 4])).astype(np.float32)
                meas[f, t] = z
                ox, oP, oSI = O.kf_update(state[t][0], state[t][1], z, conf[f, t])
                x, P = kfs[t].update(state[t][0], state[t][1], z, conf[f, t])
                np.testing.assert_allclose(ox, x, rtol=1e-11, atol=1e-9)
                np.testing.assert_allclose(oP, P, rtol=1e-9, atol=1e-9)
                np.testing.assert_allclose(oSI, kfs[t].SI, rtol=1e-9, atol=1e-12)
                state[t] = [x.copy(), P.copy(), kfs[t].SI.copy()]
        xs.append(np.stack([s[0] for s in state]))
        Ps.append(np.stack([s[1] for s in state]))
        SIs.append(np.stack([s[2] for s in state]))
    _save("kalman", z0=z0, matched=matched, meas=meas, conf=conf, x=np.stack(xs), P=np.stack(Ps), SI=np.stack(SIs))


def gold_posttricks():
    """N4: flip-TTA merge, clamp + attribute distance add, sklearn's DBSCAN on the precomputed Jaccard matrix,
    smooth_tracklets and diminish_camera_bias, each produced by the reference's own code / dependency."""
    import torch.nn.functional as TF
    from sklearn.cluster import DBSCAN
    IU = ref_import.load_file("ref_inference_utils", "reid/inference_utils.py")
    U = ref_import.load_file("ref_losses_utils", "reid/losses/utils.py")
    rng = np.random.default_rng(17)
    N, D, n_ids = 420, 32, 28
    cent = rng.standard_normal((n_ids, D)).astype(np.float32)
    ids = rng.integers(0, n_ids, N)
    spread = np.where(rng.random(N) < 0.15, 1.4, 0.35).astype(np.float32)[:, None]  # some outliers: noise / border points
    e1 = (cent[ids] + spread * rng.standard_normal((N, D))).astype(np.float32)
    e2 = (e1 + 0.05 * rng.standard_normal((N, D))).astype(np.float32)
    merged = TF.normalize((torch.from_numpy(e1) + torch.from_numpy(e2)) / 2.0, dim=1)
    np.testing.assert_allclose(O.tta_merge_normalize(e1, e2), merged.numpy(), rtol=0, atol=2e-7)
    cams = rng.integers(0, 4, N)
    unbiased = IU.diminish_camera_bias(merged.clone(), torch.from_numpy(cams).float())
    np.testing.assert_allclose(O.diminish_camera_bias(merged.numpy(), cams), unbiased.numpy(), rtol=0, atol=2e-5)
    feats = unbiased.numpy()
    J = O.compute_jaccard_distance(feats, 20, 6)  # (pinned against the reference in gold_jaccard)
    attrs = TF.normalize(torch.from_numpy(rng.integers(1, 3, (n_ids, 30)).astype(np.float32)[ids]), dim=1)
    attr_dist = U.euclidean_dist(attrs, attrs).numpy()
    out = {}
    for tag, eps, ms, other in (("plain", 0.5, 7, None), ("attr", 0.62, 5, attr_dist), ("tight", 0.3, 3, None)):
        d = J.copy()
        d[d < 0] = 0.
        if other is not None:
            d += other
        np.testing.assert_array_equal(O.clamp_add(J, other), d)
        cm = DBSCAN(eps=eps, min_samples=ms, metric="precomputed", n_jobs=-1)
        lab = cm.fit_predict(d)
        core = np.zeros(N, bool)
        core[cm.core_sample_indices_] = True
        olab, ocore = O.dbscan_precomputed(d, eps, ms)
        assert np.array_equal(olab, lab) and np.array_equal(ocore, core), tag
        border = (lab != -1) & ~core
        print("  dbscan %-5s clusters %d noise %d border %d" % (tag, lab.max() + 1, (lab == -1).sum(), border.sum()))
        assert (lab == -1).any() and lab.max() >= 3
        out["lab_" + tag], out["core_" + tag] = lab.astype(np.int64), core
        out["eps_" + tag], out["ms_" + tag] = np.float64(eps), np.int32(ms)
    assert ((out["lab_plain"] != -1) & ~out["core_plain"]).any() or ((out["lab_attr"] != -1) & ~out["core_attr"]).any()
    lab = out["lab_plain"]
    seqs = rng.integers(0, 5, N)
    num_labels = int(lab.max()) + 1
    merged_seqs = torch.from_numpy(seqs * num_labels + lab)
    valid = lab != -1
    smooth = IU.smooth_tracklets(unbiased.clone(), merged_seqs, valid)
    np.testing.assert_allclose(O.smooth_tracklets(feats, merged_seqs.numpy(), valid), smooth.numpy(), rtol=0, atol=2e-7)
    _save("posttricks", e1=e1, e2=e2, merged=merged.numpy(), cams=cams.astype(np.int32), unbiased=feats, J=J,
          attr_dist=attr_dist, merged_seqs=merged_seqs.numpy().astype(np.int64), smooth=smooth.numpy(), **out)


def main():
    assert ref_import.available(), "/root/reference not present: fixtures can only be (re)generated in the build container"
    ref_import.install()
    for fn in (gold_dist, gold_eval, gold_rerank_eval, gold_jaccard, gold_preprocess, gold_assoc, gold_boxpairs, gold_ocsort, gold_kalman, gold_posttricks):
        print(fn.__name__)
        fn()
    print("oracle pinned against the reference; fixtures in tests/golden/")


if __name__ == "__main__":
    main()
