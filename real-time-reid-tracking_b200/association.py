"""Drop-in for the OC-SORT half of yolov8_tracking/boxmot/utils/association.py: speed_direction_batch (:6-16),
linear_assignment (:18-26), associate (:107-190).  The D x T IoU matrix and the velocity-direction-consistency term
come from ONE kernel launch (csrc/boxpairs.cu: ocsort_cost_kernel); the assignment itself stays on the host (SURVEY 8f
N1) through scipy -- `lap`, which the reference prefers when installed, is not available in this image."""
from __future__ import annotations

import numpy as np

from . import ops


def speed_direction_batch(dets, tracks):
    """(dy, dx) unit directions from every track's box centre to every detection's, [T,D] (association.py:6-16)."""
    tracks = np.asarray(tracks)[..., np.newaxis]
    dets = np.asarray(dets)
    CX1, CY1 = (dets[:, 0] + dets[:, 2]) / 2.0, (dets[:, 1] + dets[:, 3]) / 2.0
    CX2, CY2 = (tracks[:, 0] + tracks[:, 2]) / 2.0, (tracks[:, 1] + tracks[:, 3]) / 2.0
    dx, dy = CX1 - CX2, CY1 - CY2
    norm = np.sqrt(dx ** 2 + dy ** 2) + 1e-6
    return dy / norm, dx / norm


def linear_assignment(cost_matrix):
    """Pairs [[row, col], ...] of the minimum-cost assignment (association.py:18-26, the lapjv branch's layout)."""
    from scipy.optimize import linear_sum_assignment

    x, y = linear_sum_assignment(cost_matrix)
    return np.array(list(zip(x, y)), dtype=int).reshape(-1, 2)


def associate_costs(detections, trackers, velocities, previous_obs, vdc_weight):
    """(iou_matrix, angle_diff_cost) numpy float64 [D,T]: the two geometric terms of associate's final_cost."""
    iou, ang = ops.ocsort_cost(detections, trackers, velocities, previous_obs, vdc_weight)
    return iou.cpu().numpy(), ang.cpu().numpy()


def associate(detections, trackers, iou_threshold, velocities, previous_obs, vdc_weight, emb_cost=None, w_assoc_emb=None,
              aw_off=None, aw_param=None):
    """(matches int[K,2], unmatched_detections, unmatched_trackers) as association.py:107-190 returns them.  Only the
    aw_off form of the embedding term is offered (compute_aw_max_metric is a host-side re-weighting, out of scope)."""
    detections, trackers = np.asarray(detections, np.float64), np.asarray(trackers, np.float64)
    if len(trackers) == 0:
        return np.empty((0, 2), dtype=int), np.arange(len(detections)), np.empty((0, 5), dtype=int)
    iou_matrix, angle_diff_cost = associate_costs(detections, trackers, velocities, previous_obs, vdc_weight)
    if min(iou_matrix.shape):
        a = (iou_matrix > iou_threshold).astype(np.int32)
        if a.sum(1).max() == 1 and a.sum(0).max() == 1:
            matched_indices = np.stack(np.where(a), axis=1)
        else:
            if emb_cost is None:
                emb = 0
            else:
                if not aw_off:
                    raise NotImplementedError("associate: adaptive embedding weighting (aw_off=False) is not provided")
                emb = np.array(emb_cost.numpy() if hasattr(emb_cost, "numpy") else emb_cost, dtype=np.float64)
                emb[iou_matrix <= 0] = 0
                emb = emb * w_assoc_emb
            matched_indices = linear_assignment(-(iou_matrix + angle_diff_cost + emb))
            if matched_indices.size == 0:
                matched_indices = np.empty(shape=(0, 2))
    else:
        matched_indices = np.empty(shape=(0, 2))
    mi = np.asarray(matched_indices).reshape(-1, 2).astype(int)
    weak = iou_matrix[mi[:, 0], mi[:, 1]] < iou_threshold  # assigned pairs that overlap too little are dissolved
    free_d = np.setdiff1d(np.arange(len(detections)), mi[:, 0])
    free_t = np.setdiff1d(np.arange(len(trackers)), mi[:, 1])
    return mi[~weak], np.concatenate([free_d, mi[weak, 0]]), np.concatenate([free_t, mi[weak, 1]])
