"""Drop-in for yolov8_tracking/boxmot/utils/matching.py.  StrongSORT half: chi2inv95 (:12-22), _pdist (:255-275),
_cosine_distance (:278-298), _nn_cosine_distance (:321-339), NearestNeighborDistanceMetric (:342-409).  ByteTrack /
BoT-SORT half: ious (:71-88), iou_distance (:91-112), embedding_distance (:138-160), fuse_iou (:189-200),
fuse_score (:203-213), bbox_ious (:216-248) on csrc/boxpairs.cu.

The per-track sample lists live in a device-resident ring buffer (ops.Gallery); `distance` is one launch of the
fused association kernel without the gate, returning the same float64 [T,M] matrix of float32 values.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib as L
from . import ops

chi2inv95 = {1: 3.8415, 2: 5.9915, 3: 7.8147, 4: 9.4877, 5: 11.070, 6: 12.592, 7: 14.067, 8: 15.507, 9: 16.919}


def _pdist(a, b):
    """Pair-wise squared distance, clipped at 0 (matching.py:255-275); float32 values from the distmat kernel."""
    a, b = np.asarray(a), np.asarray(b)
    if len(a) == 0 or len(b) == 0:
        return np.zeros((len(a), len(b)))
    r2 = ops.distmat(a, b, "sql2").cpu().numpy()
    return np.clip(r2, 0.0, float(np.inf))


def _cosine_distance(a, b, data_is_normalized=False):
    """1 - a^.b^T, rows re-normalised unless data_is_normalized (matching.py:278-298)."""
    out = ops.cosine_distance(a, b, data_is_normalized)
    return out if isinstance(a, torch.Tensor) and a.is_cuda else out.cpu().numpy()


def _nn_cosine_distance(x, y):
    """Smallest cosine distance of every row of y to a sample in x (matching.py:321-339)."""
    d = ops.cosine_distance(np.asarray(x, np.float32), np.asarray(y, np.float32))
    return d.min(dim=0)[0].cpu().numpy()


def _nn_euclidean_distance(x, y):
    """The reference's version is broken (torch.min on an ndarray); fixed by restatement."""
    return np.maximum(0.0, _pdist(x, y).min(axis=0))


def bbox_ious(boxes, query_boxes):
    """'+1' pixel IoU of float32 tlbr boxes, [N,K] float32 (matching.py:216-248; a Python double loop there)."""
    return ops.box_pairs(np.asarray(boxes, np.float32), np.asarray(query_boxes, np.float32), "plus_one").cpu().numpy()


def ious(atlbrs, btlbrs):
    """matching.py:71-88: float32 [len(a), len(b)]; empty inputs give an empty matrix without a launch."""
    out = np.zeros((len(atlbrs), len(btlbrs)), dtype=np.float32)
    if out.size == 0:
        return out
    return bbox_ious(np.ascontiguousarray(atlbrs, dtype=np.float32), np.ascontiguousarray(btlbrs, dtype=np.float32))


def _tlbrs(tracks):
    if len(tracks) > 0 and isinstance(tracks[0], np.ndarray):
        return tracks
    return [t.tlbr for t in tracks]


def iou_distance(atracks, btracks):
    """1 - IoU of the tracks' tlbr boxes (or of plain box arrays), matching.py:91-112."""
    if (len(atracks) > 0 and isinstance(atracks[0], np.ndarray)) or (len(btracks) > 0 and isinstance(btracks[0], np.ndarray)):
        atlbrs, btlbrs = atracks, btracks
    else:
        atlbrs, btlbrs = [t.tlbr for t in atracks], [t.tlbr for t in btracks]
    if len(atlbrs) == 0 or len(btlbrs) == 0:
        return 1 - np.zeros((len(atlbrs), len(btlbrs)), dtype=np.float32)
    return ops.box_pairs(np.ascontiguousarray(atlbrs, dtype=np.float32), np.ascontiguousarray(btlbrs, dtype=np.float32),
                         "iou_distance").cpu().numpy()


def embedding_distance(tracks, detections, metric="cosine"):
    """max(0, cosine distance) of the tracks' smooth_feat to the detections' curr_feat (matching.py:138-160).  The
    reference goes through scipy's float64 cdist; here the product runs on the fp32-accurate tensor-core path
    (|difference| <= 1e-6) and the matrix comes back as float64 like cdist's."""
    cost = np.zeros((len(tracks), len(detections)), dtype=np.float32)
    if cost.size == 0:
        return cost
    if metric != "cosine":
        raise ValueError("embedding_distance: only the cosine metric is used by the reference trackers")
    det = np.asarray([t.curr_feat for t in detections], dtype=np.float32)
    trk = np.asarray([t.smooth_feat for t in tracks], dtype=np.float32)
    d = ops.cosine_distance(trk, det).cpu().numpy().astype(np.float64)
    return np.maximum(0.0, d)


def fuse_iou(cost_matrix, tracks, detections):
    """1 - (1 - cost) * (1 + iou_sim) / 2 (matching.py:189-200)."""
    cost_matrix = np.asarray(cost_matrix)
    if cost_matrix.size == 0:
        return cost_matrix
    a = np.ascontiguousarray(_tlbrs(tracks), dtype=np.float32)
    b = np.ascontiguousarray(_tlbrs(detections), dtype=np.float32)
    return ops.fuse_cost(cost_matrix, tlbr_a=a, tlbr_b=b).cpu().numpy()


def fuse_score(cost_matrix, detections):
    """1 - (1 - cost) * det.score (matching.py:203-213), float64."""
    cost_matrix = np.asarray(cost_matrix)
    if cost_matrix.size == 0:
        return cost_matrix
    scores = np.array([d.score for d in detections]) if len(detections) and hasattr(detections[0], "score") else np.asarray(detections)
    return ops.fuse_cost(cost_matrix, det_scores=scores).cpu().numpy()


class NearestNeighborDistanceMetric(object):
    """Per-target nearest-neighbour appearance metric with a device-resident gallery (matching.py:342-409).

    samples: dict target -> slot; the rows themselves stay on the GPU in a [slots, ring, d] ring buffer created on the
    first partial_fit.  Like the reference's dict of lists the store is unbounded: the slot table doubles when more than
    `max_tracks` targets are alive, and with budget=None (keep every sample) the ring doubles whenever a target would
    outgrow it (ops.Gallery.resize: contents move bit for bit).  `max_tracks` / `max_unbounded` are only the initial sizes.

    metric="euclidean": the reference's helper is broken (torch.min on an ndarray raises TypeError at the first
    distance() call); here it is fixed by restatement -- the smallest SQUARED distance to a stored sample,
    clipped at 0, samples kept unnormalised -- on the distmat kernel, off the fused association path."""

    def __init__(self, metric, matching_threshold, budget=None, max_tracks=512, max_unbounded=128):
        if metric not in ("euclidean", "cosine"):
            raise ValueError("Invalid metric; must be either 'euclidean' or 'cosine'")
        self.metric_name = metric
        self.matching_threshold = matching_threshold
        self.budget = budget
        self._ring = int(budget) if budget is not None else int(max_unbounded)
        self._slots_cap = int(max_tracks)
        self._gallery = None
        self._slot_of = {}      # target id -> slot
        self._rows_of = {}      # target id -> rows held (host mirror of the device counts: no sync needed to grow)
        self._free = []
        self._raw = {}          # euclidean only: target id -> f32[n, d] device rows, oldest first

    # -- the reference exposes `samples` as dict id -> list of rows; here: id -> stored (normalised) rows, lazily
    @property
    def samples(self):
        return _SamplesView(self)

    def _ensure(self, d):
        if self._gallery is None:
            self._gallery = ops.Gallery(self._slots_cap, self._ring, d)
            self._free = list(range(self._slots_cap - 1, -1, -1))
        elif self._gallery.d != d:
            raise ValueError("feature dimension changed from %d to %d" % (self._gallery.d, d))

    def _slot(self, target):
        s = self._slot_of.get(target)
        if s is None:
            if not self._free:  # more live targets than slots: double the table (the reference's dict has no limit)
                old = self._slots_cap
                self._slots_cap = 2 * old
                self._gallery.resize(slots=self._slots_cap)
                self._free = list(range(self._slots_cap - 1, old - 1, -1))
            s = self._free.pop()
            self._slot_of[target] = s
            self._rows_of[target] = 0
        return s

    def partial_fit(self, features, targets, active_targets):
        """Append features to their targets (last `budget` survive) and forget inactive targets (matching.py:374-389)."""
        targets = [int(t) for t in np.asarray(targets).reshape(-1)]
        active = set(int(t) for t in active_targets)
        if self.metric_name == "euclidean":
            feats = L.as_f32_cuda(features) if len(targets) else None
            for i, t in enumerate(targets):
                rows = feats[i:i + 1] if t not in self._raw else torch.cat([self._raw[t], feats[i:i + 1]], 0)
                self._raw[t] = rows[-self.budget:] if self.budget is not None else rows
            self._raw = {int(k): self._raw[int(k)] for k in active_targets}  # (KeyError for an unknown target, like the reference)
            return
        if len(targets):
            feats = L.as_f32_cuda(features)
            self._ensure(feats.shape[1])
            slots = np.array([self._slot(t) for t in targets], np.int32)
            need = 0
            for t in targets:
                self._rows_of[t] += 1
                need = max(need, self._rows_of[t])
            if self.budget is None and need > self._ring:  # keep every sample: grow the ring instead of dropping rows
                while self._ring < need:
                    self._ring *= 2
                self._gallery.resize(budget=self._ring)
            elif self.budget is not None:
                for t in set(targets):
                    self._rows_of[t] = min(self._rows_of[t], self._ring)
            self._gallery.partial_fit(feats, slots)
        dead = [t for t in self._slot_of if t not in active]
        if dead:
            ds = np.array([self._slot_of.pop(t) for t in dead], np.int32)
            for t in dead:
                self._rows_of.pop(t, None)
            self._gallery.reset_slots(ds)
            self._free.extend(int(s) for s in ds)

    def slots_for(self, targets):
        return np.array([self._slot_of[int(t)] for t in np.asarray(targets).reshape(-1)], np.int32)

    def distance_device(self, features, targets):
        """f64 [T,M] CUDA tensor (no host sync)."""
        T = len(targets)
        feats = L.as_f32_cuda(features)
        M = feats.shape[0]
        if T == 0 or M == 0:
            return torch.zeros((T, M), dtype=torch.float64, device=feats.device)
        if self.metric_name == "euclidean":
            out = torch.empty((T, M), dtype=torch.float64, device=feats.device)
            for i, t in enumerate(targets):
                out[i] = ops.distmat(self._raw[int(t)], feats, "sql2").clamp_(min=0.0).min(dim=0)[0]
            return out
        slots = self.slots_for(targets)
        zeros8 = torch.zeros((T, 8), dtype=torch.float64, device=feats.device)
        tlwh = torch.ones((M, 4), dtype=torch.float32, device=feats.device)
        app, _, _ = ops.assoc_frame(self._gallery, feats, tlwh, slots, zeros8, None, np.ones(T, np.int32), gate=False,
                                    want_iou=False)
        return app

    def distance(self, features, targets):
        """np.ndarray float64 [len(targets), len(features)] (matching.py:391-409)."""
        return self.distance_device(features, targets).cpu().numpy()


class _SamplesView(object):
    def __init__(self, metric):
        self._m = metric

    def _ids(self):
        return self._m._raw if self._m.metric_name == "euclidean" else self._m._slot_of

    def __contains__(self, k):
        return int(k) in self._ids()

    def __len__(self):
        return len(self._ids())

    def keys(self):
        return self._ids().keys()

    def __iter__(self):
        return iter(self._ids())

    def __getitem__(self, k):
        """Stored rows of target k, oldest first (L2-normalised: the metric re-normalises anyway)."""
        m = self._m
        if m.metric_name == "euclidean":
            return list(m._raw[int(k)].cpu().numpy())
        slot = m._slot_of[int(k)]
        n = int(m._gallery.counts()[slot].item())
        return list(m._gallery.read_slot(slot)[:n].cpu().numpy())
