"""Per-frame association of StrongSORT's Tracker._match (yolov8_tracking/boxmot/trackers/strongsort/sort/tracker.py:175-233)
on top of ONE fused kernel launch: appearance (per-track gallery cosine minimum) + Mahalanobis gate + blend + clamp,
and the IoU cost, for every (track, detection) pair; two small float64 matrices come back for the host Hungarian solve
(scipy).  `match_arrays` works on plain arrays; `match` takes the reference's Track / Detection objects."""
from __future__ import annotations

import numpy as np
import torch

from . import ops
from .linear_assignment import _solve, track_gate_state


def frame_costs(metric, det_feat, det_tlwh, trk_ids, trk_confirmed, trk_mean, trk_SI, trk_tsu, mc_lambda, max_dist=None,
                max_iou_dist=0.7, iou_mode="iou", clamp=True, want_raw=False):
    """Device matrices (app f64[T,M], raw or None, iou f64[T,M]); rows of unconfirmed tracks carry 1e5 in `app`."""
    T = len(trk_ids)
    max_dist = metric.matching_threshold if max_dist is None else max_dist
    slots = np.full(T, -1, np.int32)
    for r in range(T):
        if trk_confirmed[r] and int(trk_ids[r]) in metric.samples:
            slots[r] = metric._slot_of[int(trk_ids[r])]
    if metric._gallery is None:  # nothing fitted yet (no confirmed track can exist): allocate an empty gallery
        metric._ensure(int(np.asarray(det_feat).shape[1]))
    gallery = metric._gallery
    # track state: host arrays, or device tensors straight from kalman.BatchedStrongSortKF.gate_state (no H2D)
    as64 = lambda a, w: a.reshape(T, w) if isinstance(a, torch.Tensor) else np.asarray(a, np.float64).reshape(T, w)
    return ops.assoc_frame(gallery, det_feat, det_tlwh, slots, as64(trk_mean, 8), as64(trk_SI, 16),
                           trk_tsu if isinstance(trk_tsu, torch.Tensor) else np.asarray(trk_tsu, np.int32), mc_lambda,
                           max_dist, max_iou_dist, iou_mode, clamp_app=clamp, clamp_iou=clamp, want_raw=want_raw)


def match_arrays(metric, det_feat, det_tlwh, trk_ids, trk_confirmed, trk_mean, trk_SI, trk_tsu, mc_lambda, max_iou_dist,
                 max_dist=None, iou_mode="iou"):
    """(matches [(track_row, det_idx)], unmatched_tracks, unmatched_detections) exactly as Tracker._match returns them."""
    T, M = len(trk_ids), len(det_tlwh)
    max_dist = metric.matching_threshold if max_dist is None else max_dist
    confirmed = [i for i in range(T) if trk_confirmed[i]]
    unconfirmed = [i for i in range(T) if not trk_confirmed[i]]
    dets = list(range(M))
    if T and M:
        app, _, iou = frame_costs(metric, det_feat, det_tlwh, trk_ids, trk_confirmed, trk_mean, trk_SI, trk_tsu, mc_lambda,
                                  max_dist, max_iou_dist, iou_mode)
        both = torch.stack([app, iou]).cpu().numpy()  # one D2H
        app, iou = both[0], both[1]
    # stage A: confirmed tracks x all detections, appearance cost (matching_cascade, single level)
    if len(confirmed) == 0 or M == 0:
        m_a, um_d = [], dets
    else:
        m_a, _, um_d = _solve(np.ascontiguousarray(app[confirmed]), max_dist, confirmed, dets)
    um_a = list(set(confirmed) - set(k for k, _ in m_a))
    # stage B: unconfirmed + just-missed confirmed tracks x remaining detections, IoU cost
    cand = unconfirmed + [k for k in um_a if trk_tsu[k] == 1]
    um_a = [k for k in um_a if trk_tsu[k] != 1]
    if len(cand) == 0 or len(um_d) == 0:
        m_b, um_b = [], cand
    else:
        m_b, um_b, um_d = _solve(np.ascontiguousarray(iou[np.ix_(cand, um_d)]), max_iou_dist, cand, um_d)
    return m_a + m_b, list(set(um_a + um_b)), um_d


def match(metric, tracks, detections, mc_lambda, max_iou_dist):
    """Tracker._match on the reference's objects (Track: track_id, is_confirmed(), time_since_update, mean, kf.SI;
    Detection: tlwh float32, feature)."""
    T, M = len(tracks), len(detections)
    mean = np.zeros((T, 8))
    SI = np.zeros((T, 16))
    for r, t in enumerate(tracks):
        mean[r] = np.asarray(t.mean, np.float64).reshape(-1)[:8]
        SI[r] = track_gate_state(t)[1].reshape(-1)
    det_tlwh = np.asarray([d.tlwh for d in detections], np.float32).reshape(M, 4)
    det_feat = np.asarray([np.asarray(d.feature) for d in detections], np.float32)
    return match_arrays(metric, det_feat, det_tlwh, [t.track_id for t in tracks], [t.is_confirmed() for t in tracks], mean, SI,
                        [t.time_since_update for t in tracks], mc_lambda, max_iou_dist)
