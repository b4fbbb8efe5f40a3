"""Drop-in for yolov8_tracking/boxmot/trackers/strongsort/sort/linear_assignment.py: min_cost_matching (:12-77),
matching_cascade (:80-144), gate_cost_matrix (:147-198).  Cost matrices come from the CUDA kernels; the Hungarian
solve (scipy.optimize.linear_sum_assignment) and the match bookkeeping stay on the host."""
from __future__ import annotations

import numpy as np
from scipy.optimize import linear_sum_assignment

from . import ops
from .matching import chi2inv95

INFTY_COST = 1e5


def min_cost_matching(distance_metric, max_distance, tracks, detections, track_indices=None, detection_indices=None):
    if track_indices is None:
        track_indices = np.arange(len(tracks))
    if detection_indices is None:
        detection_indices = np.arange(len(detections))
    if len(detection_indices) == 0 or len(track_indices) == 0:
        return [], track_indices, detection_indices  # Nothing to match.

    cost_matrix = distance_metric(tracks, detections, track_indices, detection_indices)
    cost_matrix[cost_matrix > max_distance] = max_distance + 1e-5
    return _solve(cost_matrix, max_distance, track_indices, detection_indices)


def _solve(cost_matrix, max_distance, track_indices, detection_indices):
    """Hungarian + threshold filter on an already clamped cost matrix (linear_assignment.py:60-77)."""
    row_indices, col_indices = linear_sum_assignment(cost_matrix)
    matches, unmatched_tracks, unmatched_detections = [], [], []
    for col, detection_idx in enumerate(detection_indices):
        if col not in col_indices:
            unmatched_detections.append(detection_idx)
    for row, track_idx in enumerate(track_indices):
        if row not in row_indices:
            unmatched_tracks.append(track_idx)
    for row, col in zip(row_indices, col_indices):
        track_idx = track_indices[row]
        detection_idx = detection_indices[col]
        if cost_matrix[row, col] > max_distance:
            unmatched_tracks.append(track_idx)
            unmatched_detections.append(detection_idx)
        else:
            matches.append((track_idx, detection_idx))
    return matches, unmatched_tracks, unmatched_detections


def matching_cascade(distance_metric, max_distance, cascade_depth, tracks, detections, track_indices=None,
                     detection_indices=None):
    """Single-level in this reference (the per-age loop is commented out, linear_assignment.py:129-144)."""
    if track_indices is None:
        track_indices = list(range(len(tracks)))
    if detection_indices is None:
        detection_indices = list(range(len(detections)))
    unmatched_detections = detection_indices
    matches, _, unmatched_detections = min_cost_matching(distance_metric, max_distance, tracks, detections,
                                                         track_indices, unmatched_detections)
    unmatched_tracks = list(set(track_indices) - set(k for k, _ in matches))
    return matches, unmatched_tracks, unmatched_detections


def track_gate_state(track):
    """(x[:4], SI 4x4) of a track's Kalman filter as the reference's md_for_measurement reads them (stale SI, F6)."""
    kf = getattr(track, "kf", None)
    x = np.asarray(kf.x if kf is not None and hasattr(kf, "x") else track.mean, np.float64).reshape(-1)[:4]
    SI = np.asarray(kf.SI if kf is not None and hasattr(kf, "SI") else np.zeros((4, 4)), np.float64).reshape(4, 4)
    return x, SI


def gate_cost_matrix(cost_matrix, tracks, detections, track_indices, detection_indices, mc_lambda,
                     gated_cost=INFTY_COST, only_position=False):
    """In place AND returned: cost[row, gd > chi2inv95] = gated_cost; cost[row] = l*cost[row] + (1-l)*gd."""
    gating_dim = 2 if only_position else 4
    gating_threshold = chi2inv95[gating_dim]
    T, M = len(track_indices), len(detection_indices)
    measurements = np.asarray([detections[i].to_xyah() for i in detection_indices])
    if only_position and T:
        # the reference forwards the flag to track.kf.gating_distance (linear_assignment.py:188-190), whose
        # only_position branch cannot run: reshape_z / the 2x2 product raise ValueError for every input
        # (strongsort_kf_adapter.py:160-168, "Needs to be tested!").  Same exception here, through the adapter.
        from .strongsort_kf_adapter import StrongSortKalmanFilterAdapter
        StrongSortKalmanFilterAdapter._only_position_error(measurements)
    if T == 0 or M == 0:
        return cost_matrix
    x4 = np.zeros((T, 4))
    SI = np.zeros((T, 16))
    for row, track_idx in enumerate(track_indices):
        x, s = track_gate_state(tracks[track_idx])
        x4[row], SI[row] = x, s.reshape(-1)
    gd = ops.gating_distance(x4, SI, measurements.astype(np.float64)).cpu().numpy()
    for row in range(T):
        cost_matrix[row, gd[row] > gating_threshold] = gated_cost
        cost_matrix[row] = mc_lambda * cost_matrix[row] + (1 - mc_lambda) * gd[row]
    return cost_matrix
