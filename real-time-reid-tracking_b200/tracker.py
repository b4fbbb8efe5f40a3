"""Per-frame association of StrongSORT's Tracker._match (yolov8_tracking/boxmot/trackers/strongsort/sort/tracker.py:175-233)
on top of ONE fused kernel launch: appearance (per-track gallery cosine minimum) + Mahalanobis gate + blend + clamp,
and the IoU cost, for every (track, detection) pair; two small float64 matrices come back for the host Hungarian solve
(scipy).  `match_arrays` works on plain arrays; `match` takes the reference's Track / Detection objects."""
from __future__ import annotations

import numpy as np
import torch

from . import _lib as L
from . import ops
from .linear_assignment import _solve, track_gate_state


def frame_costs(metric, det_feat, det_tlwh, trk_ids, trk_confirmed, trk_mean, trk_SI, trk_tsu, mc_lambda, max_dist=None,
                max_iou_dist=0.7, iou_mode="iou", clamp=True, want_raw=False):
    """Device matrices (app f64[T,M], raw or None, iou f64[T,M]); rows of unconfirmed tracks carry 1e5 in `app`."""
    T = len(trk_ids)
    max_dist = metric.matching_threshold if max_dist is None else max_dist
    slot_of = metric._slot_of  # id -> ring-buffer slot of every target that has samples
    slots = np.fromiter((slot_of.get(int(i), -1) if c else -1 for i, c in zip(trk_ids, trk_confirmed)), np.int32, T)
    if metric._gallery is None:  # nothing fitted yet (no confirmed track can exist): allocate an empty gallery
        metric._ensure(int(np.asarray(det_feat).shape[1]))
    gallery = metric._gallery
    dev = gallery.device
    # per-frame host state goes up in ONE pinned copy; device tensors (kalman.BatchedStrongSortKF.gate_state) pass through
    host = {}
    if not isinstance(trk_mean, torch.Tensor):
        host["mean"] = np.ascontiguousarray(np.asarray(trk_mean, np.float64).reshape(T, 8))
    if not isinstance(trk_SI, torch.Tensor):
        host["SI"] = np.ascontiguousarray(np.asarray(trk_SI, np.float64).reshape(T, 16))
    if not isinstance(det_tlwh, torch.Tensor):
        host["tlwh"] = np.ascontiguousarray(np.asarray(det_tlwh, np.float32).reshape(-1, 4))
    if not isinstance(trk_tsu, torch.Tensor):
        host["tsu"] = np.ascontiguousarray(np.asarray(trk_tsu, np.int32).reshape(T))
    host["slots"] = slots
    up = dict(zip(host.keys(), L.staging(dev).upload(list(host.values()))))
    mean_d = up.get("mean", trk_mean.reshape(T, 8) if isinstance(trk_mean, torch.Tensor) else None)
    SI_d = up.get("SI", trk_SI.reshape(T, 16) if isinstance(trk_SI, torch.Tensor) else None)
    M = int(up["tlwh"].shape[0]) if "tlwh" in up else int(det_tlwh.shape[0])
    out = torch.empty((2, T, M), dtype=torch.float64, device=dev)  # app | iou in one buffer: one D2H for the host solve
    return ops.assoc_frame(gallery, det_feat, up.get("tlwh", det_tlwh), up["slots"], mean_d, SI_d, up.get("tsu", trk_tsu),
                           mc_lambda, max_dist, max_iou_dist, iou_mode, clamp_app=clamp, clamp_iou=clamp, want_raw=want_raw,
                           out_app=out[0], out_iou=out[1])


class FrameContext:
    """A tracker loop's per-frame host <-> device traffic through two C calls (csrc/assoc.cu reidb200_frame_*): `crops`
    before the backbone, `costs` after it.  Host arrays go in as plain pointers (they must already be C-contiguous with the
    dtypes below: no per-frame conversion), each call makes ONE pinned host->device copy; `costs` returns numpy views of
    the pinned result of ONE device->host copy.  One context per video stream (it owns the staging buffers).

        fc = FrameContext(metric, max_dets=256, max_tracks=512)
        batch = fc.crops(frame_cuda_u8, xyxy_int32)                 # f32[M,3,256,128] on the device, asynchronous
        feats = backbone(batch)                                     # f32[M,d] CUDA
        app, iou = fc.costs(feats, det_tlwh_f32, trk_ids, trk_confirmed, mean_f64, SI_f64, tsu_i32, 0.995, 0.2, 0.7)

    app / iou are float64 [T,M] exactly as tracker.frame_costs(...clamp=True) + costs_to_host return them; they stay valid
    until the call after the next."""

    def __init__(self, metric, max_dets=256, max_tracks=512, out_size=(128, 256), norm=((0.5, 0.5, 0.5), (0.5, 0.5, 0.5))):
        import ctypes as C

        self.metric = metric
        self.max_dets, self.max_tracks = int(max_dets), int(max_tracks)
        self.out_w, self.out_h = int(out_size[0]), int(out_size[1])
        self._mean, self._std = ops._f3(norm[0]), ops._f3(norm[1])
        self._lib = L.load()
        self._h = None
        self._gallery = None
        self._C = C
        self._pa, self._pi = C.c_void_p(), C.c_void_p()
        self._crops = None

    def _handle(self, d=None):
        if self._h is None or self._gallery is not self.metric._gallery:
            if self.metric._gallery is None:
                self.metric._ensure(int(d))
            self.close()
            self._gallery = self.metric._gallery
            h = self._C.c_void_p()
            with torch.cuda.device(self._gallery.device):
                L.check(self._lib.reidb200_frame_create(self._gallery._h, self.max_dets, self.max_tracks, self._C.byref(h)))
            self._h = h
        return self._h

    def close(self):
        if getattr(self, "_h", None) is not None and L is not None:
            self._lib.reidb200_frame_destroy(self._h)
            self._h = None

    __del__ = close

    def crops(self, frame, xyxy, out=None):
        """frame: u8[H,W,3] CUDA tensor; xyxy: int32 [M,4] host array -> f32[M,3,out_h,out_w] CUDA (asynchronous)."""
        n = int(xyxy.shape[0])
        H, W = int(frame.shape[0]), int(frame.shape[1])
        if out is None:
            if self._crops is None or self._crops.shape[0] < n:
                self._crops = torch.empty((max(n, self.max_dets), 3, self.out_h, self.out_w), dtype=torch.float32, device=frame.device)
            out = self._crops[:n]
        if n:
            if xyxy.dtype != np.int32 or not xyxy.flags.c_contiguous:
                xyxy = np.ascontiguousarray(xyxy, dtype=np.int32)
            L.check(self._lib.reidb200_frame_crops(self._handle(), frame.data_ptr(), H, W, xyxy.ctypes.data, n, self.out_h, self.out_w,
                                                   self._mean, self._std, out.data_ptr(), L.stream_ptr()))
        return out

    def costs(self, det_feat, det_tlwh, trk_ids, trk_confirmed, trk_mean, trk_SI, trk_tsu, mc_lambda, max_dist=None,
              max_iou_dist=0.7, iou_mode="iou", clamp=True, slots=None):
        """(app f64[T,M], iou f64[T,M]) numpy views of the pinned result.  det_feat: f32[M,d] CUDA; det_tlwh f32[M,4],
        trk_mean f64[T,8], trk_SI f64[T,16] (or [T,4,4]), trk_tsu i32[T]: C-contiguous host arrays.  `slots` (i32[T],
        -1 = no gallery row, e.g. unconfirmed tracks) may be passed to skip the id -> slot lookup."""
        T, M = len(trk_tsu), int(det_tlwh.shape[0])
        if T == 0 or M == 0:
            z = np.zeros((T, M))
            return z, z.copy()
        if slots is None:
            slot_of = self.metric._slot_of
            slots = np.fromiter((slot_of.get(int(i), -1) if c else -1 for i, c in zip(trk_ids, trk_confirmed)), np.int32, T)
        max_dist = self.metric.matching_threshold if max_dist is None else max_dist
        flags = (L.ASSOC_CLAMP_APP | L.ASSOC_CLAMP_IOU) if clamp else 0
        f32, f64, i32 = np.float32, np.float64, np.int32
        det_tlwh = det_tlwh if det_tlwh.dtype == f32 and det_tlwh.flags.c_contiguous else np.ascontiguousarray(det_tlwh, f32)
        trk_mean = trk_mean if trk_mean.dtype == f64 and trk_mean.flags.c_contiguous else np.ascontiguousarray(trk_mean, f64)
        trk_SI = trk_SI if trk_SI.dtype == f64 and trk_SI.flags.c_contiguous else np.ascontiguousarray(trk_SI, f64)
        trk_tsu = trk_tsu if isinstance(trk_tsu, np.ndarray) and trk_tsu.dtype == i32 and trk_tsu.flags.c_contiguous else np.ascontiguousarray(trk_tsu, i32)
        L.check(self._lib.reidb200_frame_costs(self._handle(det_feat.shape[1]), det_feat.data_ptr(), det_feat.stride(0), det_tlwh.ctypes.data, M,
                                               slots.ctypes.data, trk_mean.ctypes.data, trk_SI.ctypes.data, trk_tsu.ctypes.data, T,
                                               float(mc_lambda), float(max_dist), float(max_iou_dist),
                                               L.IOU_DIOU if iou_mode == "diou" else L.IOU_PLAIN, flags,
                                               self._C.byref(self._pa), self._C.byref(self._pi), L.stream_ptr()))
        app = np.ctypeslib.as_array(self._C.cast(self._pa, self._C.POINTER(self._C.c_double)), shape=(T, M))
        iou = np.ctypeslib.as_array(self._C.cast(self._pi, self._C.POINTER(self._C.c_double)), shape=(T, M))
        return app, iou


_host_out = {"bufs": [None, None], "turn": 0}


def costs_to_host(app, iou):
    """(app, iou) numpy float64 with ONE device->host copy into pinned memory when both are the halves of
    frame_costs' buffer.  The arrays are views of one of two alternating pinned buffers: they stay valid until the
    call after the next (a frame's costs are consumed by that frame's assignment)."""
    base = app._base
    if (base is not None and base is iou._base and base.dim() == 3 and base.shape[0] == 2
            and app.data_ptr() == base.data_ptr() and iou.data_ptr() == base.data_ptr() + app.numel() * 8):
        k = _host_out["turn"]
        _host_out["turn"] = k ^ 1
        buf = _host_out["bufs"][k]
        if buf is None or buf.numel() < base.numel():
            buf = torch.empty(max(base.numel(), 1 << 16), dtype=torch.float64).pin_memory()
            _host_out["bufs"][k] = buf
        h = buf[:base.numel()].view(base.shape)
        h.copy_(base, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        h = h.numpy()
        return h[0], h[1]
    return app.cpu().numpy(), iou.cpu().numpy()


def match_arrays(metric, det_feat, det_tlwh, trk_ids, trk_confirmed, trk_mean, trk_SI, trk_tsu, mc_lambda, max_iou_dist,
                 max_dist=None, iou_mode="iou", device_lap=True):
    """(matches [(track_row, det_idx)], unmatched_tracks, unmatched_detections) exactly as Tracker._match returns them.
    device_lap=True (the default since the exact mode exists) solves both assignment stages on the device too (SURVEY 8f
    N1) with scipy's own assignment, ties included: the three lists come back in the reference's order (lap.cu exact
    mode).  device_lap=False runs scipy on the host, the reference's own solver.  device_lap="fast" is the
    component-parallel solver: the same matches whenever scipy's optimum is unique, unmatched detections in ascending
    order (scipy's pairing of gated leftovers, which orders the reference's list, is not reproduced)."""
    T, M = len(trk_ids), len(det_tlwh)
    max_dist = metric.matching_threshold if max_dist is None else max_dist
    confirmed = [i for i in range(T) if trk_confirmed[i]]
    unconfirmed = [i for i in range(T) if not trk_confirmed[i]]
    dets = list(range(M))
    if T and M:
        app, _, iou = frame_costs(metric, det_feat, det_tlwh, trk_ids, trk_confirmed, trk_mean, trk_SI, trk_tsu, mc_lambda,
                                  max_dist, max_iou_dist, iou_mode)
        if device_lap == "fast":
            return _match_on_device(app, iou, confirmed, unconfirmed, trk_tsu, max_dist, max_iou_dist)
        if device_lap:
            return _match_on_device_exact(app, iou, confirmed, unconfirmed, trk_tsu, max_dist, max_iou_dist)
        app, iou = costs_to_host(app, iou)  # one D2H
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


_lap_host = {}


def _match_on_device(app, iou, confirmed, unconfirmed, trk_tsu, max_dist, max_iou_dist):
    """Both stages of Tracker._match (tracker.py:175-233) chained on the device: stage A over the confirmed rows of the
    appearance matrix, stage B over the IoU matrix for the unconfirmed rows and the confirmed rows with
    time_since_update == 1 that stage A left unmatched (the kernel filters rows / columns matched so far).  One small
    H2D (the two row lists) and one D2H (2T + M + 1 ints) per frame."""
    dev = app.device
    T, M = app.shape
    tsu = np.asarray(trk_tsu)
    conf_a = np.asarray(confirmed, np.int32)
    cand_b = np.concatenate([np.asarray(unconfirmed, np.int32), conf_a[tsu[conf_a] == 1] if conf_a.size else conf_a]).astype(np.int32)
    init = _lap_init.get((T, M))
    if init is None:
        _lap_init.clear()
        init = _lap_init[(T, M)] = np.concatenate([np.full(T, -1, np.int32), np.zeros(M + 1 + T, np.int32)])
    # row lists and the initial state (match = -1, det_matched = 0, status = 0, match after stage A) in one H2D
    rows_a, rows_b, state = L.staging(dev).upload([conf_a, cand_b, init])
    match, det_m, status, match_a = state[:T], state[T:T + M], state[T + M:T + M + 1], state[T + M + 1:]
    if conf_a.size:
        ops.lap_gated(app, rows_a, max_dist, match, det_m, status)
    match_a.copy_(match)
    if cand_b.size:
        ops.lap_gated(iou, rows_b, max_iou_dist, match, det_m, status)
    key = (dev.index, state.numel())
    host = _lap_host.get(key)
    if host is None:
        _lap_host.clear()
        host = _lap_host[key] = torch.empty(state.numel(), dtype=torch.int32).pin_memory()
    host.copy_(state, non_blocking=True)
    torch.cuda.current_stream(dev).synchronize()
    h = host.numpy()
    if h[T + M]:
        raise L.Reidb200Error("device assignment met non-finite costs")
    mt, mt_a = h[:T], h[T + M + 1:]
    ra = conf_a[mt_a[conf_a] >= 0] if conf_a.size else conf_a
    rb = cand_b[(mt[cand_b] >= 0) & (mt_a[cand_b] < 0)] if cand_b.size else cand_b
    matches = list(zip(ra.tolist(), mt_a[ra].tolist())) + list(zip(rb.tolist(), mt[rb].tolist()))
    return matches, np.nonzero(mt < 0)[0].tolist(), np.nonzero(h[T:T + M] == 0)[0].tolist()


def _match_on_device_exact(app, iou, confirmed, unconfirmed, trk_tsu, max_dist, max_iou_dist):
    """Both stages of Tracker._match (tracker.py:175-233) on the device with the reference's list orders: stage A =
    min_cost_matching of the confirmed rows x all detections on the appearance matrix, stage B = the unconfirmed rows +
    the confirmed rows with time_since_update == 1 that stage A left unmatched x stage A's unmatched_detections IN ITS
    ORDER on the IoU matrix.  The lists between and after the stages are built by the reference's own expressions
    (list(set(...)) included: CPython's set order decides the row order of stage B), so stage A's row -> column table
    comes back to the host before stage B is launched; stage A's detection list stays on the device."""
    dev = app.device
    T, M = app.shape
    tsu = np.asarray(trk_tsu)
    conf_a = np.asarray(confirmed, np.int32)
    # state: match[T] = -1 | um_a[M] | n_a[1] | det_matched[M] = 0 | status[1] | um_b[M] | n_b[1] | um_rows[T] | n_rows[1]
    init = _lap_init.get(("x", T, M))
    if init is None:
        init = _lap_init[("x", T, M)] = np.concatenate([np.full(T, -1, np.int32), np.zeros(3 * (M + 1) + T + 1, np.int32)])
    dets = _lap_init.get(("d", M))
    if dets is None:
        dets = _lap_init[("d", M)] = np.arange(M, dtype=np.int32)
    stg = L.staging(dev)
    rows_a, cols_a, state = stg.upload([conf_a, dets, init])
    o = 0
    match = state[o:o + T]; o += T
    um_a = state[o:o + M]; o += M
    n_a = state[o:o + 1]; o += 1
    det_m = state[o:o + M]; o += M
    status = state[o:o + 1]; o += 1
    um_b = state[o:o + M]; o += M
    n_b = state[o:o + 1]; o += 1
    um_rows = state[o:o + T]; o += T
    n_rows = state[o:o + 1]
    host = _lap_host.get((dev.index, state.numel()))
    if host is None:
        host = _lap_host[(dev.index, state.numel())] = torch.empty(state.numel(), dtype=torch.int32).pin_memory()
    h = host.numpy()
    stream = torch.cuda.current_stream(dev)
    # ---- stage A
    if conf_a.size:
        ops.lap_exact(app, rows_a, cols_a, None, max_dist, match, det_m, um_a, n_a, status)
        host.copy_(state, non_blocking=True)  # (a few KB: the whole state, so that the status word is current too)
        stream.synchronize()
        mt_a = h[:T].copy()
        ra = conf_a[mt_a[conf_a] >= 0]
        m_a = list(zip(ra.tolist(), mt_a[ra].tolist()))
        n_um = int(h[T + M])
    else:
        um_a.copy_(cols_a)
        n_a.fill_(M)
        m_a, n_um = [], M
    um_trk_a = list(set(confirmed) - set(ra.tolist())) if conf_a.size else []
    # ---- stage B
    cand = list(unconfirmed) + [k for k in um_trk_a if tsu[k] == 1]
    um_trk_a = [k for k in um_trk_a if tsu[k] != 1]
    if len(cand) == 0 or n_um == 0:
        m_b, um_trk_b = [], cand
        if not conf_a.size:
            host.copy_(state, non_blocking=True)
            stream.synchronize()
        um_d = h[T:T + n_um].tolist()
    else:
        cand_np = np.asarray(cand, np.int32)
        (rows_b,) = stg.upload([cand_np])
        ops.lap_exact(iou, rows_b, um_a, n_a, max_iou_dist, match, det_m, um_b, n_b, status, um_rows, n_rows)
        host.copy_(state, non_blocking=True)
        stream.synchronize()
        mt = h[:T]
        rb = cand_np[mt[cand_np] >= 0]
        m_b = list(zip(rb.tolist(), mt[rb].tolist()))
        ob = T + 2 * (M + 1)
        um_d = h[ob:ob + int(h[ob + M])].tolist()
        orr = ob + M + 1
        um_trk_b = h[orr:orr + int(h[orr + T])].tolist()
    if h[T + 2 * M + 1]:
        raise L.Reidb200Error("device assignment met non-finite costs")
    return m_a + m_b, list(set(um_trk_a + um_trk_b)), um_d


_lap_init = {}


def match(metric, tracks, detections, mc_lambda, max_iou_dist, device_lap=True):
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
                        [t.time_since_update for t in tracks], mc_lambda, max_iou_dist, device_lap=device_lap)
