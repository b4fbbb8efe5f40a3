"""Batched, device-resident form of StrongSORT's Kalman filter.

Reference: one StrongSortKalmanFilterAdapter object per Track (boxmot/trackers/strongsort/sort/track.py:97-98), each
predict / update a handful of 8 x 8 numpy products from a Python loop over the tracks (sort/tracker.py:64-70, 106-140).
Here the states of all tracks live in three device arrays indexed by slot, a frame's predictions are ONE launch and a
frame's updates ONE launch (csrc/kalman.cu), and `mean` / `SI` feed ops.assoc_frame without a host round trip."""
from __future__ import annotations

import numpy as np
import torch

from . import _lib as L


class BatchedStrongSortKF:
    std_weight_position = 1.0 / 20
    std_weight_velocity = 1.0 / 160

    def __init__(self, max_tracks=1024, dt=1.0, device=None):
        self.dev = device or L.require_device()
        self.dt = float(dt)
        self.mean = torch.zeros((max_tracks, 8), dtype=torch.float64, device=self.dev)
        self.covariance = torch.zeros((max_tracks, 8, 8), dtype=torch.float64, device=self.dev)
        self.SI = torch.zeros((max_tracks, 16), dtype=torch.float64, device=self.dev)

    def _slots(self, slots):
        s = torch.as_tensor(np.ascontiguousarray(np.asarray(slots, dtype=np.int32)) if not isinstance(slots, torch.Tensor) else slots)
        s = s.to(device=self.dev, dtype=torch.int32).reshape(-1).contiguous()
        if s.numel() and (int(s.max()) >= self.mean.shape[0] or int(s.min()) < 0):
            raise IndexError("track slot out of range")
        return s

    def _f64(self, a, width):
        t = torch.as_tensor(np.ascontiguousarray(np.asarray(a, dtype=np.float64)) if not isinstance(a, torch.Tensor) else a)
        return t.to(device=self.dev, dtype=torch.float64).reshape(-1, width).contiguous()

    def initiate(self, slots, measurements):
        """New tracks from xyah measurements [n,4] (strongsort_kf_adapter.py:25-58)."""
        s, m = self._slots(slots), self._f64(measurements, 4)
        if m.shape[0] != s.numel():
            raise ValueError("one measurement per slot expected")
        with torch.cuda.device(self.dev):
            L.check(L.load().reidb200_kf_initiate(self.mean.data_ptr(), self.covariance.data_ptr(), self.SI.data_ptr(),
                                                 s.data_ptr(), m.data_ptr(), s.numel(), self.std_weight_position,
                                                 self.std_weight_velocity, L.stream_ptr()))

    def predict(self, slots):
        """One prediction step for the listed tracks (Tracker.predict's loop, sort/tracker.py:64-70)."""
        s = self._slots(slots)
        with torch.cuda.device(self.dev):
            L.check(L.load().reidb200_kf_predict(self.mean.data_ptr(), self.covariance.data_ptr(), s.data_ptr(), s.numel(),
                                                self.dt, self.std_weight_position, self.std_weight_velocity, L.stream_ptr()))

    def update(self, slots, measurements, confidences=None):
        """Measurement update of the matched tracks (Track.update -> kf.update, sort/track.py:176-181)."""
        s, m = self._slots(slots), self._f64(measurements, 4)
        if m.shape[0] != s.numel():
            raise ValueError("one measurement per slot expected")
        c = None
        if confidences is not None:
            c = self._f64(confidences, 1).reshape(-1)
            if c.numel() != s.numel():
                raise ValueError("one confidence per slot expected")
        with torch.cuda.device(self.dev):
            L.check(L.load().reidb200_kf_update(self.mean.data_ptr(), self.covariance.data_ptr(), self.SI.data_ptr(),
                                               s.data_ptr(), m.data_ptr(), L.ptr(c), s.numel(), self.std_weight_position,
                                               L.stream_ptr()))

    def gate_state(self, slots):
        """(mean f64[n,8], SI f64[n,16]) device tensors in the order of `slots`: the trk_mean / trk_SI arguments of
        ops.assoc_frame / tracker.frame_costs."""
        idx = self._slots(slots).long()
        return self.mean.index_select(0, idx), self.SI.index_select(0, idx)
