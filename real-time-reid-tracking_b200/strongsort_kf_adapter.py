"""Drop-in for yolov8_tracking/boxmot/motion/adapters/strongsort_kf_adapter.py: StrongSortKalmanFilterAdapter with the
reference's per-track object interface (initiate :25-58, predict :60-94, update :96-135, gating_distance :137-169) on
the device kernels of csrc/kalman.cu / csrc/assoc.cu.

One object per Track, as boxmot/trackers/strongsort/sort/track.py:97-98 creates them.  The state lives in a slot of a
shared kalman.BatchedStrongSortKF pool on the device; `x`, `P`, `SI` are host mirrors refreshed by every call (the
reference's attributes).  The per-object calls exist for drop-in use; the batched pool (one launch per frame for all
tracks) is the fast path and what tracker.frame_costs consumes without a host round trip."""
from __future__ import annotations

import numpy as np
import torch

from . import _lib as L
from . import ops
from .kalman import BatchedStrongSortKF


class _Pool:
    """Slots of one BatchedStrongSortKF per device, handed out to adapter objects and recycled when they die."""

    def __init__(self, device, capacity=1024):
        self.kf = BatchedStrongSortKF(capacity, device=device)
        self.free = list(range(capacity - 1, -1, -1))

    def take(self):
        if not self.free:
            old = self.kf.mean.shape[0]
            grown = BatchedStrongSortKF(2 * old, dt=self.kf.dt, device=self.kf.dev)
            grown.mean[:old].copy_(self.kf.mean)
            grown.covariance[:old].copy_(self.kf.covariance)
            grown.SI[:old].copy_(self.kf.SI)
            self.kf = grown
            self.free = list(range(2 * old - 1, old - 1, -1))
        return self.free.pop()


_pools = {}


def _pool(device):
    key = (device.type, device.index)
    if key not in _pools:
        _pools[key] = _Pool(device)
    return _pools[key]


class StrongSortKalmanFilterAdapter(object):
    ndim = 4

    def __init__(self, dt=1):
        self._dev = L.require_device()
        self._pool = _pool(self._dev)
        self._slot = self._pool.take()
        self._dt = float(dt)
        self.dim_x, self.dim_z = 8, 4
        self._std_weight_position = 1.0 / 20
        self._std_weight_velocity = 1.0 / 160
        self.F = np.eye(8)
        for i in range(self.ndim):
            self.F[i, self.ndim + i] = dt
        self.H = np.eye(self.ndim, 2 * self.ndim)
        self.x = np.zeros(8)
        self.P = np.eye(8)
        self.SI = np.zeros((4, 4))  # inverse innovation covariance as of the last update (stale afterwards)

    def __del__(self):
        pool = getattr(self, "_pool", None)
        if pool is not None and getattr(self, "_slot", None) is not None:
            pool.free.append(self._slot)
            self._slot = None

    # ---- state plumbing
    def _kf(self):
        kf = self._pool.kf
        kf.dt = self._dt
        kf.std_weight_position = self._std_weight_position
        kf.std_weight_velocity = self._std_weight_velocity
        return kf

    def _push(self, mean, covariance):
        kf, s = self._kf(), self._slot
        kf.mean[s].copy_(torch.as_tensor(np.asarray(mean, np.float64).reshape(8)))
        kf.covariance[s].copy_(torch.as_tensor(np.asarray(covariance, np.float64).reshape(8, 8)))

    def _pull(self):
        kf, s = self._kf(), self._slot
        self.x = kf.mean[s].cpu().numpy()
        self.P = kf.covariance[s].cpu().numpy()
        self.SI = kf.SI[s].cpu().numpy().reshape(4, 4)

    # ---- the reference's interface
    def initiate(self, measurement):
        """(mean f64[8], covariance f64[8,8]) of a new track from an xyah measurement; SI starts at zero."""
        self._kf().initiate([self._slot], np.asarray(measurement, np.float64).reshape(1, 4))
        self._pull()
        return self.x, self.P

    def predict(self, mean, covariance):
        """One prediction step; like the reference it advances the filter's OWN state (self.x / self.P set by the
        previous call) -- the arguments only supply the values the motion noise is scaled by, which are the same."""
        self._kf().predict([self._slot])
        self._pull()
        return self.x, self.P

    def update(self, mean, covariance, measurement, confidence=0.0):
        self._push(mean, covariance)
        self._kf().update([self._slot], np.asarray(measurement, np.float64).reshape(1, 4), [float(confidence)])
        self._pull()
        return self.x, self.P

    @staticmethod
    def _only_position_error(measurements):
        """The reference's only_position branch (:160-168) cannot run: it hands the whole N x 4 array to reshape_z
        (ValueError unless N == 1) and then multiplies a (4,2)-broadcast residual with a 2 x 2 block (ValueError)."""
        n = int(np.asarray(measurements).shape[0])
        if n != 1:
            raise ValueError("z must be convertible to shape (4, 1)")
        raise ValueError("shapes (2,4) and (2,2) not aligned: 4 (dim 1) != 2 (dim 0)")

    def gating_distance(self, measurements, only_position=False):
        """f64[N]: sqrt(y^T SI y), y = z - H x, per measurement (xyah rows), with the SI stored by the last update --
        what KalmanFilter.md_for_measurement returns (kalman_filter.py:577-586); one kernel launch for all N."""
        measurements = np.asarray(measurements)
        if measurements.shape[0] == 0:
            return np.zeros((0,))
        if only_position:
            self._only_position_error(measurements)
        gd = ops.gating_distance(np.asarray(self.x, np.float64)[:4], np.asarray(self.SI, np.float64).reshape(16),
                                 measurements.astype(np.float64))
        return gd.cpu().numpy()[0]
