"""Drop-in for yolov8_tracking/boxmot/trackers/strongsort/sort/iou_matching.py (iou :9-45, iou_cost :48-86) and
the DIoU variant of modification_deepsort/iou_matching.py (:5-47)."""
from __future__ import annotations

import numpy as np

from . import ops

INFTY_COST = 1e5


def iou(bbox, candidates):
    """Plain IoU of one tlwh box against M tlwh candidates -> float64 [M]."""
    candidates = np.asarray(candidates)
    if candidates.size == 0:
        return np.zeros((0,))
    return ops.iou_matrix(np.asarray(bbox, np.float64).reshape(1, 4), candidates.reshape(-1, 4), "iou").cpu().numpy()[0]


def diou(bbox, candidates):
    """modification_deepsort/iou_matching.py::iou: IoU - centre distance^2 / outer diagonal^2."""
    candidates = np.asarray(candidates)
    if candidates.size == 0:
        return np.zeros((0,))
    return ops.iou_matrix(np.asarray(bbox, np.float64).reshape(1, 4), candidates.reshape(-1, 4), "diou").cpu().numpy()[0]


def iou_cost(tracks, detections, track_indices=None, detection_indices=None, use_diou=False):
    """float64 [len(track_indices), len(detection_indices)]: 1 - iou, rows with time_since_update > 1 = 1e5."""
    if track_indices is None:
        track_indices = np.arange(len(tracks))
    if detection_indices is None:
        detection_indices = np.arange(len(detections))
    T, M = len(track_indices), len(detection_indices)
    if T == 0 or M == 0:
        return np.zeros((T, M))
    mean = np.zeros((T, 8))
    tsu = np.zeros(T, np.int32)
    for r, ti in enumerate(track_indices):
        tr = tracks[ti]
        tl = np.asarray(tr.to_tlwh(), np.float64)  # exact inverse below: xyah such that to_tlwh reproduces tl
        mean[r, :4] = np.asarray(tr.mean, np.float64).reshape(-1)[:4] if hasattr(tr, "mean") else _tlwh_to_xyah(tl)
        tsu[r] = tr.time_since_update
    cands = np.asarray([detections[i].tlwh for i in detection_indices], np.float32)
    _, _, c = ops.assoc_frame(None, None, cands, None, mean, None, tsu, iou_mode="diou" if use_diou else "iou",
                              want_app=False)
    return c.cpu().numpy()


def _tlwh_to_xyah(tl):
    r = np.array(tl, np.float64)
    r[:2] += r[2:] / 2
    r[2] /= r[3]
    return r
