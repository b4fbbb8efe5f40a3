"""Tensor-level bindings: torch CUDA tensors in, torch CUDA tensors out, every call one or more sm_100a
kernels of libreidb200.so on the current torch stream.  The reference-named modules (evaluate.py,
losses_utils.py, faiss_utils.py, matching.py, ...) are thin shims over these."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib as L

_KIND = {"euclid": L.DIST_EUCLID, "cosine_half": L.DIST_COSINE_HALF, "dot": L.DIST_DOT, "sql2": L.DIST_SQL2}


def _f32(x, device=None):
    return L.as_f32_cuda(x, device)


def _i32(x, device):
    if not isinstance(x, torch.Tensor):
        x = torch.from_numpy(np.ascontiguousarray(np.asarray(x)))
    return x.to(device=device, dtype=torch.int32, non_blocking=True).contiguous()


def _f3(v):
    return (C.c_float * 3)(*[float(a) for a in v])


# ----------------------------------------------------------------------------------------- distances
def distmat(x, y, kind="euclid", out=None):
    x = _f32(x)
    dev = L.require_device(x)
    y = _f32(y, dev)
    m, d = x.shape
    n, d2 = y.shape
    if d != d2:
        raise ValueError("feature dimensions differ: %d vs %d" % (d, d2))
    if out is None:
        out = torch.empty((m, n), dtype=torch.float32, device=dev)
    if m == 0 or n == 0:
        return out
    lib = L.load()
    ws = L.workspace(lib.reidb200_distmat_workspace_bytes(m, n, d), dev)
    with torch.cuda.device(dev):
        L.check(lib.reidb200_distmat(x.data_ptr(), m, x.stride(0), y.data_ptr(), n, y.stride(0), d, _KIND[kind],
                                     out.data_ptr(), out.stride(0), ws.data_ptr(), ws.numel(), L.stream_ptr()))
    return out


def distmat_simt(x, y, kind="euclid"):
    """fp32 SIMT cross-check kernel (tests only)."""
    x = _f32(x)
    dev = L.require_device(x)
    y = _f32(y, dev)
    m, d = x.shape
    n = y.shape[0]
    out = torch.empty((m, n), dtype=torch.float32, device=dev)
    if m and n:
        with torch.cuda.device(dev):
            L.check(L.load().reidb200_distmat_simt(x.data_ptr(), m, x.stride(0), y.data_ptr(), n, y.stride(0), d,
                                                  _KIND[kind], out.data_ptr(), out.stride(0), L.stream_ptr()))
    return out


def knn(xq, xb, k, metric="sql2", index_base=0, idx64=True):
    """Exact brute-force k-NN: (D f32[nq,k], I i64|i32[nq,k]), ascending for sql2, descending for dot."""
    xq = _f32(xq)
    dev = L.require_device(xq)
    xb = _f32(xb, dev)
    nq, d = xq.shape
    nb = xb.shape[0]
    D = torch.empty((nq, k), dtype=torch.float32, device=dev)
    I = torch.empty((nq, k), dtype=torch.int64 if idx64 else torch.int32, device=dev)
    if nq == 0:
        return D, I
    if nb == 0:
        raise ValueError("knn: empty database")
    lib = L.load()
    ws = L.workspace(lib.reidb200_knn_workspace_bytes(nq, nb, d, k), dev)
    with torch.cuda.device(dev):
        L.check(lib.reidb200_knn(xq.data_ptr(), nq, xq.stride(0), xb.data_ptr(), nb, xb.stride(0), d, k, _KIND[metric],
                                 int(index_base), D.data_ptr(), I.data_ptr(), 1 if idx64 else 0, ws.data_ptr(),
                                 ws.numel(), L.stream_ptr()))
    return D, I


def pack_rows(x, out=None, sqnorm=None, iscale=None):
    """f32[rows,d] -> (packed u8 bytes of the scaled fp16 hi|lo planes, sqnorm f32[rows], iscale f32[rows]): the resident
    form of a gallery.  `out` / `sqnorm` / `iscale` may be slices of a larger preallocated shard (piecewise packing)."""
    x = _f32(x)
    dev = L.require_device(x)
    rows, d = x.shape
    lib = L.load()
    nbytes = lib.reidb200_pack_bytes(rows, d)
    if out is None:
        out = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=dev)
    if sqnorm is None:
        sqnorm = torch.empty(rows, dtype=torch.float32, device=dev)
    if iscale is None:
        iscale = torch.empty(rows, dtype=torch.float32, device=dev)
    if rows:
        with torch.cuda.device(dev):
            L.check(lib.reidb200_pack_rows(x.data_ptr(), rows, x.stride(0), d, out.data_ptr(), sqnorm.data_ptr(),
                                           iscale.data_ptr(), L.stream_ptr()))
    return out, sqnorm, iscale


def packed_row_bytes(d):
    """bytes one gallery row occupies in packed form (fp16 hi + lo planes, d padded to 64)."""
    return 2 * ((int(d) + 63) // 64 * 64) * 2


def knn_packed(xq, packed, sqnorm, iscale, nb, d, k, metric="sql2", index_base=0, idx64=True):
    """knn against a gallery packed once by pack_rows (resident index / one shard of a sharded gallery)."""
    xq = _f32(xq)
    dev = L.require_device(xq)
    nq, dq = xq.shape
    if dq != d:
        raise ValueError("knn_packed: query dimension %d != gallery dimension %d" % (dq, d))
    D = torch.empty((nq, k), dtype=torch.float32, device=dev)
    I = torch.empty((nq, k), dtype=torch.int64 if idx64 else torch.int32, device=dev)
    if nq == 0:
        return D, I
    if nb == 0:
        raise ValueError("knn: empty database")
    lib = L.load()
    ws = L.workspace(lib.reidb200_knn_packed_workspace_bytes(nq, nb, d, k), dev)
    with torch.cuda.device(dev):
        L.check(lib.reidb200_knn_packed(xq.data_ptr(), nq, xq.stride(0), packed.data_ptr(), sqnorm.data_ptr(), iscale.data_ptr(), int(nb),
                                        int(d),
                                        int(k), _KIND[metric], int(index_base), D.data_ptr(), I.data_ptr(), 1 if idx64 else 0,
                                        ws.data_ptr(), ws.numel(), L.stream_ptr()))
    return D, I


def topk_rows(vals, k, descending=False, cand_i=None, index_base=0, idx64=True):
    vals = _f32(vals)
    dev = L.require_device(vals)
    rows, cols = vals.shape
    D = torch.empty((rows, k), dtype=torch.float32, device=dev)
    I = torch.empty((rows, k), dtype=torch.int64 if idx64 else torch.int32, device=dev)
    if cand_i is not None:
        cand_i = cand_i.to(device=dev, dtype=torch.int64).contiguous()
        assert cand_i.shape == vals.shape
    if rows:
        with torch.cuda.device(dev):
            L.check(L.load().reidb200_topk_rows(vals.data_ptr(), rows, cols, vals.stride(0), L.ptr(cand_i), k,
                                               1 if descending else 0, int(index_base), D.data_ptr(), I.data_ptr(),
                                               1 if idx64 else 0, L.stream_ptr()))
    return D, I


# ----------------------------------------------------------------------------------------- CMC / mAP
def eval_cmc_map(mat, ql, qc, gl, gc, descending=True, per_query=False):
    """mat f32[Q,G] on device -> (cmc f32[G], ap f64[1]) device tensors (+ ap_q f64[Q], first i32[Q])."""
    mat = _f32(mat)
    dev = L.require_device(mat)
    Q, G = mat.shape
    ql, qc, gl, gc = _i32(ql, dev), _i32(qc, dev), _i32(gl, dev), _i32(gc, dev)
    if ql.numel() != Q or qc.numel() != Q or gl.numel() != G or gc.numel() != G:
        raise ValueError("label / camera vectors do not match the score matrix")
    cmc = torch.empty(G, dtype=torch.float32, device=dev)
    ap = torch.empty(1, dtype=torch.float64, device=dev)
    ap_q = torch.empty(Q, dtype=torch.float64, device=dev) if per_query else None
    first = torch.empty(Q, dtype=torch.int32, device=dev) if per_query else None
    lib = L.load()
    ws = L.workspace(lib.reidb200_eval_workspace_bytes(Q, G), dev)
    with torch.cuda.device(dev):
        L.check(lib.reidb200_eval_cmc_map(mat.data_ptr(), Q, G, mat.stride(0), 1 if descending else 0, ql.data_ptr(),
                                          qc.data_ptr(), gl.data_ptr(), gc.data_ptr(), cmc.data_ptr(), ap.data_ptr(),
                                          L.ptr(ap_q), L.ptr(first), ws.data_ptr(), ws.numel(), L.stream_ptr()))
    if per_query:
        return cmc, ap, ap_q, first
    return cmc, ap


# ----------------------------------------------------------------------------------------- re-ranking
def jaccard(feats, k1=20, k2=6, use_float16=False, lambda_value=0.0, rows=None, cols=None, want_rank=False):
    """Jaccard distance of the merged set `feats` f32[N,d]; rows/cols = (start, count) sub-block.
    Returns the device matrix (f32 or f16) and optionally the initial ranking i32[N,k1]."""
    feats = _f32(feats)
    dev = L.require_device(feats)
    N, d = feats.shape
    row0, nrows = rows if rows is not None else (0, N)
    col0, ncols = cols if cols is not None else (0, N)
    out = torch.empty((nrows, ncols), dtype=torch.float16 if use_float16 else torch.float32, device=dev)
    rank = torch.empty((N, k1), dtype=torch.int32, device=dev) if want_rank else None
    lib = L.load()
    ws = L.workspace(lib.reidb200_jaccard_region_workspace_bytes(N, d, k1, k2, nrows if lambda_value > 0 else 0), dev)
    with torch.cuda.device(dev):
        L.check(lib.reidb200_jaccard(feats.data_ptr(), N, feats.stride(0), d, k1, k2, 1 if use_float16 else 0,
                                     float(lambda_value), row0, nrows, col0, ncols, out.data_ptr(), out.stride(0),
                                     L.ptr(rank), ws.data_ptr(), ws.numel(), L.stream_ptr()))
    return (out, rank) if want_rank else out


def rerank_rows_from_rank(feats, rank_i, rank_d, k1, k2, rows, cols, lambda_value=0.0, blend_sql2=None, use_float16=False,
                          want_out=True, labels=None):
    """Stages D3-D8 for the output block rows x cols ((start, count) each) of one merged set from a GIVEN global initial
    ranking (rank_i i32[N,k1], rank_d f32[N,k1] squared L2): what one rank of a multi-GPU re-ranking runs after the k-NN
    tables have been gathered (csrc/rerank.cu reidb200_rerank_rows_from_rank).  blend_sql2 f32[nrows, >= ncols]: squared L2
    of the block, needed when lambda_value > 0.  labels = (ql i32[nrows], qc, gl i32[ncols], gc) adds the fused evaluation:
    returns (out or None, hist i32[ncols], ap_q f64[nrows], first i32[nrows]); else returns out."""
    feats = _f32(feats)
    dev = L.require_device(feats)
    N, d = feats.shape
    rank_i = _i32(rank_i, dev)
    rank_d = _f32(rank_d, dev)
    if rank_i.shape != (N, k1) or rank_d.shape != (N, k1):
        raise ValueError("rerank_rows_from_rank: the ranking tables must be [N, k1]")
    row0, nrows = rows
    col0, ncols = cols
    out = torch.empty((nrows, ncols), dtype=torch.float16 if use_float16 else torch.float32, device=dev) if want_out else None
    ldb = 0
    if lambda_value > 0:
        if blend_sql2 is None or blend_sql2.shape[0] != nrows or blend_sql2.shape[1] < ncols or blend_sql2.stride(1) != 1:
            raise ValueError("rerank_rows_from_rank: blend_sql2 f32[nrows, >= ncols] is required when lambda_value > 0")
        ldb = blend_sql2.stride(0)
    hist = ap_q = first = status = None
    lab = [None] * 4
    if labels is not None:
        lab = [_i32(x, dev) for x in labels]
        if lab[0].numel() != nrows or lab[1].numel() != nrows or lab[2].numel() != ncols or lab[3].numel() != ncols:
            raise ValueError("rerank_rows_from_rank: label / camera vectors do not match the block")
        hist = torch.zeros(ncols, dtype=torch.int32, device=dev)
        ap_q = torch.zeros(nrows, dtype=torch.float64, device=dev)
        first = torch.full((nrows,), -1, dtype=torch.int32, device=dev)
        status = torch.zeros(64, dtype=torch.int32, device=dev)
    lib = L.load()
    ws = L.workspace(lib.reidb200_rerank_rows_workspace_bytes(N, d, k1, k2), dev)
    with torch.cuda.device(dev):
        L.check(lib.reidb200_rerank_rows_from_rank(feats.data_ptr(), N, feats.stride(0), d, rank_i.data_ptr(), rank_d.data_ptr(), k1, k2,
                                                   1 if use_float16 else 0, float(lambda_value), row0, nrows, col0, ncols,
                                                   L.ptr(blend_sql2) if lambda_value > 0 else 0, ldb, L.ptr(out), ncols,
                                                   L.ptr(lab[0]), L.ptr(lab[1]), L.ptr(lab[2]), L.ptr(lab[3]), L.ptr(hist), L.ptr(ap_q),
                                                   L.ptr(first), L.ptr(status), ws.data_ptr(), ws.numel(), L.stream_ptr()))
    if labels is not None:
        global _last_status
        _last_status = status  # (status words of the last fused evaluation: overflow flag + the kernels' debug counters)
        return out, hist, ap_q, first
    return out


_last_status = None


def k_reciprocal_lists(rank, k):
    """R(i,k) for every row: (idx i32[N,k+1] padded with -1, cnt i32[N])."""
    dev = L.require_device(rank if isinstance(rank, torch.Tensor) else None)
    rank = _i32(rank, dev)
    N, k1c = rank.shape
    idx = torch.empty((N, k + 1), dtype=torch.int32, device=dev)
    cnt = torch.empty(N, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        L.check(L.load().reidb200_k_reciprocal(rank.data_ptr(), N, k1c, k, idx.data_ptr(), cnt.data_ptr(), L.stream_ptr()))
    return idx, cnt


# ----------------------------------------------------------------------------------------- crop / resize
def crop_resize_norm(src, crop_off, crop_geom, out_h=256, out_w=128, mean=(0.5, 0.5, 0.5), std=(0.5, 0.5, 0.5),
                     mode="cv_linear", out=None):
    """src u8 device bytes, crop_off i64[M], crop_geom i32[M,3]=(h,w,stride) -> f32[M,3,out_h,out_w]."""
    dev = L.require_device(src)
    M = crop_off.numel()
    if out is None:
        out = torch.empty((M, 3, out_h, out_w), dtype=torch.float32, device=dev)
    if M == 0:
        return out
    lib = L.load()
    with torch.cuda.device(dev):
        if mode == "cv_linear":
            L.check(lib.reidb200_crop_resize_norm(src.data_ptr(), crop_off.data_ptr(), crop_geom.data_ptr(), M, out_h, out_w,
                                                  _f3(mean), _f3(std), L.RESIZE_CV_LINEAR, out.data_ptr(), L.stream_ptr()))
        elif mode == "pil_antialias":
            max_h = int(crop_geom[:, 0].max().item()) if crop_geom.is_cuda else int(crop_geom[:, 0].max())
            ws = L.workspace(lib.reidb200_resize_pil_workspace_bytes(M, max_h, out_w), dev)
            L.check(lib.reidb200_crop_resize_norm_ws(src.data_ptr(), crop_off.data_ptr(), crop_geom.data_ptr(), M, out_h,
                                                     out_w, _f3(mean), _f3(std), L.RESIZE_PIL_ANTIALIAS, out.data_ptr(),
                                                     ws.data_ptr(), ws.numel(), L.stream_ptr()))
        else:
            raise ValueError("unknown resize mode %r" % (mode,))
    return out


def resize_pil_u8(src, crop_off, crop_geom, max_h, out_h=256, out_w=128):
    dev = L.require_device(src)
    M = crop_off.numel()
    out = torch.empty((M, out_h, out_w, 3), dtype=torch.uint8, device=dev)
    if M == 0:
        return out
    lib = L.load()
    ws = L.workspace(lib.reidb200_resize_pil_workspace_bytes(M, int(max_h), out_w), dev)
    with torch.cuda.device(dev):
        L.check(lib.reidb200_resize_pil_u8(src.data_ptr(), crop_off.data_ptr(), crop_geom.data_ptr(), M, out_h, out_w,
                                           out.data_ptr(), ws.data_ptr(), ws.numel(), L.stream_ptr()))
    return out


RERANK_EVAL_MAX_G = 16384  # gallery columns one CTA of the fused kernel keeps in registers (JR_PF * JR_THREADS)


def rerank_eval(merged, G, ql, qc, gl, gc, k1=20, k2=6, lambda_value=0.3, want_dist=False, per_query=False):
    """Fused re-ranked evaluation of merged = [gallery (G rows); query]: (cmc f32[G], ap f64[1][, dist f32[Q,G]][, ap_q
    f64[Q], first i32[Q]]) on device.  Same values as jaccard(rows=(G,Q), cols=(0,G), lambda) followed by
    eval_cmc_map(descending=False); the re-ranked matrix is only materialised when want_dist; per_query adds what the
    reference's evaluate() returns for every query (AP, first-hit rank with -1 = no valid gallery match)."""
    merged = _f32(merged)
    dev = L.require_device(merged)
    N, d = merged.shape
    Q = N - int(G)
    if Q <= 0 or G <= 0:
        raise ValueError("rerank_eval: empty query or gallery set")
    if G > RERANK_EVAL_MAX_G:
        dist = jaccard(merged, k1, k2, False, lambda_value, rows=(G, Q), cols=(0, G))
        res = eval_cmc_map(dist, ql, qc, gl, gc, descending=False, per_query=per_query)
        return (res[0], res[1]) + ((dist,) if want_dist else ()) + tuple(res[2:])
    ql, qc, gl, gc = _i32(ql, dev), _i32(qc, dev), _i32(gl, dev), _i32(gc, dev)
    if ql.numel() != Q or qc.numel() != Q or gl.numel() != G or gc.numel() != G:
        raise ValueError("label / camera vectors do not match the query / gallery sets")
    lib = L.load()
    cmc = torch.empty(G, dtype=torch.float32, device=dev)
    ap = torch.empty(1, dtype=torch.float64, device=dev)
    dist = torch.empty((Q, G), dtype=torch.float32, device=dev) if want_dist else None
    ap_q = torch.empty(Q, dtype=torch.float64, device=dev) if per_query else None
    first = torch.empty(Q, dtype=torch.int32, device=dev) if per_query else None
    ws = L.workspace(lib.reidb200_rerank_eval_workspace_bytes(N, d, k1, k2, Q, G), dev)
    with torch.cuda.device(dev):
        L.check(lib.reidb200_rerank_eval_pq(merged.data_ptr(), int(G), Q, merged.stride(0), d, k1, k2, float(lambda_value), ql.data_ptr(),
                                            qc.data_ptr(), gl.data_ptr(), gc.data_ptr(), L.ptr(dist), G, cmc.data_ptr(), ap.data_ptr(),
                                            L.ptr(ap_q), L.ptr(first), ws.data_ptr(), ws.numel(), L.stream_ptr()))
    return (cmc, ap) + ((dist,) if want_dist else ()) + ((ap_q, first) if per_query else ())


class RerankEvalPlan:
    """One fused re-ranked evaluation at a fixed shape, captured once as a CUDA graph and replayed: the ~12 kernel
    launches and memsets of reidb200_rerank_eval cost one graph launch, and every buffer (features, labels, workspace,
    results) is owned by the plan.  Fill `merged` ([gallery; query] rows) and `ql/qc/gl/gc` (views of one int32 buffer),
    then run(); `cmc` / `ap` hold the result of the latest replay on the stream run() was called on.  Same values as
    rerank_eval on the same inputs (the same kernels run)."""

    def __init__(self, G, Q, D, k1=20, k2=6, lambda_value=0.3, device=None):
        G, Q, D = int(G), int(Q), int(D)
        if Q <= 0 or G <= 0:
            raise ValueError("RerankEvalPlan: empty query or gallery set")
        if G > RERANK_EVAL_MAX_G:
            raise ValueError("RerankEvalPlan: G = %d exceeds the fused kernel's %d gallery columns" % (G, RERANK_EVAL_MAX_G))
        self.dev = dev = device or L.require_device()
        self.G, self.Q, self.D, self.k1, self.k2, self.lambda_value = G, Q, D, int(k1), int(k2), float(lambda_value)
        lib = L.load()
        N = G + Q
        with torch.cuda.device(dev):
            self.merged = torch.empty((N, D), dtype=torch.float32, device=dev).normal_()
            self.lab = torch.zeros(2 * N, dtype=torch.int32, device=dev)
            self.ql, self.qc, self.gl, self.gc = self.lab[:Q], self.lab[Q:2 * Q], self.lab[2 * Q:2 * Q + G], self.lab[2 * Q + G:]
            self.cmc = torch.empty(G, dtype=torch.float32, device=dev)
            self.ap = torch.empty(1, dtype=torch.float64, device=dev)
            self._ws = L.workspace(lib.reidb200_rerank_eval_workspace_bytes(N, D, self.k1, self.k2, Q, G), dev)
            # first call outside the capture (one-time kernel attributes), on well-formed random rows
            side = torch.cuda.Stream(dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                self._launch()
            side.synchronize()
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph, stream=side):
                self._launch()

    def _launch(self):
        L.check(L.load().reidb200_rerank_eval(self.merged.data_ptr(), self.G, self.Q, self.merged.stride(0), self.D, self.k1,
                                              self.k2, self.lambda_value, self.ql.data_ptr(), self.qc.data_ptr(),
                                              self.gl.data_ptr(), self.gc.data_ptr(), 0, self.G, self.cmc.data_ptr(),
                                              self.ap.data_ptr(), self._ws.data_ptr(), self._ws.numel(), L.stream_ptr()))

    def run(self):
        self.graph.replay()
        return self.cmc, self.ap


# ----------------------------------------------------------------------------------------- association
class Gallery:
    """Device-resident per-track feature ring buffer [slots, budget, d] (NearestNeighborDistanceMetric.samples)."""

    def __init__(self, slots, budget, d, device=None):
        self.device = device or L.require_device()
        self.slots, self.budget, self.d = int(slots), int(budget), int(d)
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            L.check(L.load().reidb200_gallery_create(self.slots, self.budget, self.d, C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None) and L is not None:  # (module globals are gone at interpreter shutdown)
            L.load().reidb200_gallery_destroy(self._h)
            self._h = None

    __del__ = close

    def resize(self, slots=None, budget=None):
        """Grow in place (more slots and / or a longer ring per slot); stored rows keep their bits and their order."""
        slots = self.slots if slots is None else int(slots)
        budget = self.budget if budget is None else int(budget)
        with torch.cuda.device(self.device):
            L.check(L.load().reidb200_gallery_resize(self._h, slots, budget, L.stream_ptr()))
        self.slots, self.budget = slots, budget

    def partial_fit(self, feats, slot_of_row):
        feats = _f32(feats, self.device)
        slot_of_row = _i32(slot_of_row, self.device)
        n = feats.shape[0]
        if n:
            with torch.cuda.device(self.device):
                L.check(L.load().reidb200_gallery_partial_fit(self._h, feats.data_ptr(), feats.stride(0),
                                                             slot_of_row.data_ptr(), n, L.stream_ptr()))

    def reset_slots(self, slots):
        slots = _i32(slots, self.device)
        if slots.numel():
            with torch.cuda.device(self.device):
                L.check(L.load().reidb200_gallery_reset_slots(self._h, slots.data_ptr(), slots.numel(), L.stream_ptr()))

    def counts(self):
        out = torch.empty(self.slots, dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            L.check(L.load().reidb200_gallery_counts(self._h, out.data_ptr(), L.stream_ptr()))
        return out

    def read_slot(self, slot):
        out = torch.empty((self.budget, self.d), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            L.check(L.load().reidb200_gallery_read_slot(self._h, int(slot), out.data_ptr(), L.stream_ptr()))
        return out


def assoc_frame(gallery, det_feat, det_tlwh, trk_slot, trk_mean, trk_SI, trk_tsu, mc_lambda=0.995, max_dist=0.2,
                max_iou_dist=0.7, iou_mode="iou", clamp_app=False, clamp_iou=False, gate=True, want_raw=False,
                want_app=True, want_iou=True, simt=False, out_app=None, out_iou=None):
    """One fused launch (simt=True: the fp32 SIMT cross-check kernel instead of the tcgen05 one): (app f64[T,M] or None, raw or None, iou f64[T,M] or None), all on device."""
    dev = gallery.device if gallery is not None else L.require_device()
    det_tlwh = _f32(det_tlwh, dev)
    M = det_tlwh.shape[0]
    trk_mean = torch.as_tensor(trk_mean).to(device=dev, dtype=torch.float64).contiguous()
    T = trk_mean.shape[0]
    trk_tsu = _i32(trk_tsu, dev)
    app = raw = iou = None
    det_feat_t = None
    slot_t = si_t = None
    if want_app:
        det_feat_t = _f32(det_feat, dev)
        slot_t = _i32(trk_slot, dev)
        si_t = None if trk_SI is None else torch.as_tensor(trk_SI).to(device=dev, dtype=torch.float64).contiguous()  # (None: gate=False)
        app = out_app if out_app is not None else torch.empty((T, M), dtype=torch.float64, device=dev)
        raw = torch.empty((T, M), dtype=torch.float64, device=dev) if want_raw else None
    if want_iou:
        iou = out_iou if out_iou is not None else torch.empty((T, M), dtype=torch.float64, device=dev)
    if T == 0 or M == 0:
        return app, raw, iou
    flags = (L.ASSOC_CLAMP_APP if clamp_app else 0) | (L.ASSOC_CLAMP_IOU if clamp_iou else 0) | (0 if gate else L.ASSOC_NO_GATE) | (L.ASSOC_SIMT if simt else 0)
    with torch.cuda.device(dev):
        L.check(L.load().reidb200_assoc_frame(gallery._h if gallery is not None else None, L.ptr(det_feat_t),
                                             det_feat_t.stride(0) if det_feat_t is not None else 0, det_tlwh.data_ptr(), M,
                                             L.ptr(slot_t), trk_mean.data_ptr(), L.ptr(si_t), trk_tsu.data_ptr(), T,
                                             float(mc_lambda), float(max_dist), float(max_iou_dist),
                                             L.IOU_DIOU if iou_mode == "diou" else L.IOU_PLAIN, flags, L.ptr(app), L.ptr(raw),
                                             L.ptr(iou), L.stream_ptr()))
    return app, raw, iou


def iou_matrix(bbox_tlwh, cand_tlwh, mode="iou"):
    """T boxes f64 tlwh vs M candidates (float32 or float64 arithmetic as numpy would) -> f64[T,M] device."""
    dev = L.require_device()
    b = torch.as_tensor(np.asarray(bbox_tlwh, dtype=np.float64)).reshape(-1, 4).to(dev)
    c_np = np.asarray(cand_tlwh)
    is64 = c_np.dtype != np.float32
    c = torch.as_tensor(np.ascontiguousarray(c_np, dtype=np.float64 if is64 else np.float32)).reshape(-1, 4).to(dev)
    out = torch.empty((b.shape[0], c.shape[0]), dtype=torch.float64, device=dev)
    if b.shape[0] and c.shape[0]:
        with torch.cuda.device(dev):
            L.check(L.load().reidb200_iou_matrix(b.data_ptr(), b.shape[0], c.data_ptr(), 1 if is64 else 0, c.shape[0],
                                                L.IOU_DIOU if mode == "diou" else L.IOU_PLAIN, out.data_ptr(), L.stream_ptr()))
    return out


def _boxes(x, dtype, dev):
    t = torch.as_tensor(np.ascontiguousarray(np.asarray(x.detach().cpu() if isinstance(x, torch.Tensor) else x, dtype=dtype)))
    return t.reshape(-1, 4).to(dev)


def box_pairs(boxes_a, boxes_b, mode="iou", dtype=None):
    """Pairwise xyxy box metric [N,K] on device: iou / giou / diou / ciou (input dtype, float32 or float64), plus_one
    (bbox_ious) and iou_distance (float32).  boxmot/utils/iou.py:4-164, boxmot/utils/matching.py:71-114,216-248."""
    dev = L.require_device()
    if mode in ("plus_one", "iou_distance"):
        dtype = np.float32
    elif dtype is None:
        dtype = np.float32 if np.asarray(boxes_a).dtype == np.float32 and np.asarray(boxes_b).dtype == np.float32 else np.float64
    a, b = _boxes(boxes_a, dtype, dev), _boxes(boxes_b, dtype, dev)
    out = torch.empty((a.shape[0], b.shape[0]), dtype=torch.float32 if dtype == np.float32 else torch.float64, device=dev)
    if a.shape[0] and b.shape[0]:
        with torch.cuda.device(dev):
            L.check(L.load().reidb200_box_pairs(a.data_ptr(), a.shape[0], b.data_ptr(), b.shape[0], 0 if dtype == np.float32 else 1,
                                               L.BOX_MODES[mode], out.data_ptr(), L.stream_ptr()))
    return out


def fuse_cost(cost, det_scores=None, tlbr_a=None, tlbr_b=None):
    """fuse_score (det_scores given -> f64) or fuse_iou (both box sets given -> cost's dtype); matching.py:189-213."""
    dev = L.require_device()
    c = torch.as_tensor(cost)
    if c.dtype not in (torch.float32, torch.float64):
        c = c.to(torch.float64)
    c = c.to(dev).contiguous()
    n, k = c.shape
    is64 = c.dtype == torch.float64
    sc = a = b = None
    if det_scores is not None:
        sc = torch.as_tensor(np.asarray(det_scores, dtype=np.float64)).reshape(-1).to(dev)
        if sc.numel() != k:
            raise ValueError("fuse_score: one score per detection (column) expected")
        out = torch.empty((n, k), dtype=torch.float64, device=dev)
    else:
        a, b = _boxes(tlbr_a, np.float32, dev), _boxes(tlbr_b, np.float32, dev)
        if a.shape[0] != n or b.shape[0] != k:
            raise ValueError("fuse_iou: box sets do not match the cost matrix")
        out = torch.empty((n, k), dtype=c.dtype, device=dev)
    if n and k:
        with torch.cuda.device(dev):
            L.check(L.load().reidb200_fuse_cost(c.data_ptr(), 1 if is64 else 0, n, k, L.ptr(sc), L.ptr(a), L.ptr(b), out.data_ptr(),
                                               L.stream_ptr()))
    return out


def ocsort_cost(detections, trackers, velocities, previous_obs, vdc_weight):
    """(iou f64[D,T], angle_diff_cost f64[D,T]) device tensors of OC-SORT's associate (association.py:107-140)."""
    dev = L.require_device()
    f = lambda x, w: torch.as_tensor(np.ascontiguousarray(np.asarray(x, dtype=np.float64))).reshape(-1, w).to(dev)
    det, trk, vel, prev = f(detections, 5), f(trackers, 5), f(velocities, 2), f(previous_obs, 5)
    D, T = det.shape[0], trk.shape[0]
    if vel.shape[0] != T or prev.shape[0] != T:
        raise ValueError("ocsort_cost: one velocity and one previous observation per tracker expected")
    iou = torch.empty((D, T), dtype=torch.float64, device=dev)
    ang = torch.empty((D, T), dtype=torch.float64, device=dev)
    if D and T:
        with torch.cuda.device(dev):
            L.check(L.load().reidb200_ocsort_cost(det.data_ptr(), D, trk.data_ptr(), T, vel.data_ptr(), prev.data_ptr(),
                                                 float(vdc_weight), iou.data_ptr(), ang.data_ptr(), L.stream_ptr()))
    return iou, ang


def gating_distance(x4, SI, meas):
    dev = L.require_device()
    x4 = torch.as_tensor(np.asarray(x4, dtype=np.float64)).reshape(-1, 4).to(dev)
    SI = torch.as_tensor(np.asarray(SI, dtype=np.float64)).reshape(-1, 16).to(dev)
    meas = torch.as_tensor(np.asarray(meas, dtype=np.float64)).reshape(-1, 4).to(dev)
    out = torch.empty((x4.shape[0], meas.shape[0]), dtype=torch.float64, device=dev)
    if x4.shape[0] and meas.shape[0]:
        with torch.cuda.device(dev):
            L.check(L.load().reidb200_gating_distance(x4.data_ptr(), SI.data_ptr(), meas.data_ptr(), x4.shape[0],
                                                     meas.shape[0], out.data_ptr(), L.stream_ptr()))
    return out


def cosine_distance(a, b, data_is_normalized=False):
    a = _f32(a)
    dev = L.require_device(a)
    b = _f32(b, dev)
    na, d = a.shape
    nb = b.shape[0]
    out = torch.empty((na, nb), dtype=torch.float32, device=dev)
    if na and nb:
        lib = L.load()
        ws = L.workspace(lib.reidb200_cosine_distance_workspace_bytes(na, nb, d), dev)
        with torch.cuda.device(dev):
            L.check(lib.reidb200_cosine_distance(a.data_ptr(), na, a.stride(0), b.data_ptr(), nb, b.stride(0), d,
                                                 1 if data_is_normalized else 0, out.data_ptr(), ws.data_ptr(), ws.numel(),
                                                 L.stream_ptr()))
    return out


def jaccard_from_sql2(sql2, k1=20, k2=6, use_float16=False, lambda_value=0.0, rows=None, cols=None, sqnorm=None,
                      want_rank=False):
    """Stages D2(select)-D7 from a materialised squared-L2 matrix f32[N,N] of the merged set.
    sqnorm defaults to ones (L2-normalised features, where 2 - 2 x.y == sql2)."""
    sql2 = _f32(sql2)
    dev = L.require_device(sql2)
    N = sql2.shape[0]
    if sql2.shape[1] != N:
        raise ValueError("jaccard_from_sql2: square matrix expected")
    if sqnorm is None:
        sqnorm = torch.ones(N, dtype=torch.float32, device=dev)
    sqnorm = _f32(sqnorm, dev)
    _, rank = topk_rows(sql2, k1, descending=False, idx64=False)
    row0, nrows = rows if rows is not None else (0, N)
    col0, ncols = cols if cols is not None else (0, N)
    out = torch.empty((nrows, ncols), dtype=torch.float16 if use_float16 else torch.float32, device=dev)
    lib = L.load()
    d_any = 64  # feature dimension only sizes the dense part of the workspace, unused on this path
    ws = L.workspace(lib.reidb200_jaccard_workspace_bytes(N, d_any, k1, k2), dev)
    with torch.cuda.device(dev):
        L.check(lib.reidb200_jaccard_from_rank(rank.data_ptr(), sql2.data_ptr(), sql2.stride(0), sqnorm.data_ptr(), N, k1, k2,
                                               1 if use_float16 else 0, float(lambda_value), row0, nrows, col0, ncols,
                                               out.data_ptr(), out.stride(0), ws.data_ptr(), ws.numel(), L.stream_ptr()))
    return (out, rank) if want_rank else out


# ----------------------------------------------------------------------------------------- N4: around the Jaccard matrix
def tta_merge_normalize(a, b=None, out=None):
    """F.normalize((a + b) / 2, dim=1) (b None: plain row normalisation), f32[rows, D] on device."""
    a = _f32(a)
    dev = L.require_device(a)
    b = _f32(b, dev) if b is not None else None
    if b is not None and b.shape != a.shape:
        raise ValueError("tta_merge_normalize: shapes differ")
    rows, D = a.shape
    if out is None:
        out = torch.empty((rows, D), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        L.check(L.load().reidb200_tta_merge_normalize(a.data_ptr(), a.stride(0), L.ptr(b), b.stride(0) if b is not None else 0,
                                                      rows, D, out.data_ptr(), out.stride(0), L.stream_ptr()))
    return out


def _rows_f32_cuda(t, what):
    if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == torch.float32 and t.dim() == 2 and t.stride(1) == 1):
        raise ValueError("%s: a float32 CUDA matrix with unit column stride is required (it is modified / read in place)" % what)
    return t


def clamp_add_(dist, other=None):
    """In place on the device matrix: dist[dist < 0] = 0, then dist += other (f32[rows, cols], optional)."""
    _rows_f32_cuda(dist, "clamp_add_")
    dev = dist.device
    if other is not None:
        other = _f32(other, dev)
        if other.shape != dist.shape:
            raise ValueError("clamp_add_: shapes differ")
    rows, cols = dist.shape
    with torch.cuda.device(dev):
        L.check(L.load().reidb200_clamp_add(dist.data_ptr(), rows, cols, dist.stride(0), L.ptr(other),
                                            other.stride(0) if other is not None else 0, L.stream_ptr()))
    return dist


def dbscan_precomputed(dist, eps, min_samples):
    """sklearn's DBSCAN on a SYMMETRIC precomputed f32[N, N] device matrix (rows may be padded: stride(0) >= N; the union
    pass reads the lower triangle only, so d[i, j] must equal d[j, i] -- cluster.DBSCAN.fit checks a sample of pairs):
    (labels i32[N] with -1 = noise, n_clusters i32[1], core i32[N] 0/1), all on device, no host sync."""
    _rows_f32_cuda(dist, "dbscan_precomputed")
    N = dist.shape[0]
    if dist.shape[1] != N:
        raise ValueError("dbscan_precomputed: the distance matrix must be square")
    dev = dist.device
    labels = torch.empty(N, dtype=torch.int32, device=dev)
    ncl = torch.zeros(1, dtype=torch.int32, device=dev)
    core = torch.empty(N, dtype=torch.int32, device=dev)
    if N:
        lib = L.load()
        ws = L.workspace(lib.reidb200_dbscan_workspace_bytes(N), dev)
        with torch.cuda.device(dev):
            L.check(lib.reidb200_dbscan_precomputed(dist.data_ptr(), N, dist.stride(0), float(eps), int(min_samples), labels.data_ptr(),
                                                    ncl.data_ptr(), core.data_ptr(), ws.data_ptr(), ws.numel(), L.stream_ptr()))
    return labels, ncl, core


def smooth_groups_(emb, order, seg_start, keep=0.1, avg_w=0.9):
    """In place: rows order[seg_start[g] : seg_start[g+1]] of emb become emb * keep + mean(those rows) * avg_w."""
    _rows_f32_cuda(emb, "smooth_groups_")
    dev = emb.device
    order, seg_start = _i32(order, dev), _i32(seg_start, dev)
    n_groups = seg_start.numel() - 1
    if n_groups > 0:
        with torch.cuda.device(dev):
            L.check(L.load().reidb200_smooth_groups(emb.data_ptr(), emb.shape[1], emb.stride(0), order.data_ptr(), seg_start.data_ptr(),
                                                   n_groups, float(keep), float(avg_w), L.stream_ptr()))
    return emb


def center_columns(x, want_mean=False):
    """x - x.mean(dim=0) (f32[rows, D] on device); optionally also the mean."""
    x = _f32(x)
    dev = L.require_device(x)
    rows, D = x.shape
    out = torch.empty((rows, D), dtype=torch.float32, device=dev)
    mean = torch.empty(D, dtype=torch.float32, device=dev) if want_mean else None
    with torch.cuda.device(dev):
        L.check(L.load().reidb200_center_columns(x.data_ptr(), rows, D, x.stride(0), out.data_ptr(), out.stride(0), L.ptr(mean),
                                                 L.stream_ptr()))
    return (out, mean) if want_mean else out


# ----------------------------------------------------------------------------------------- N1: assignment on device
def lap_exact(cost, row_cand, col_cand, col_count, max_distance, match_col_of_row, det_matched, um_dets, um_count, status=None,
              um_rows=None, um_rows_count=None):
    """min_cost_matching (linear_assignment.py:52-77) on the device cost matrix f64[T, M] with scipy's OWN assignment, ties
    included (csrc/lap.cu lap_exact_kernel): rows `row_cand` (i32 device, the reference's track_indices order) x columns
    `col_cand[:col_count]` (i32 device list + i32[1] device count, or col_count=None for the whole list; the reference's
    detection_indices order); rows / columns matched so far are skipped.  Matches are recorded in match_col_of_row i32[T] /
    det_matched i32[M]; um_dets i32[>= len(col_cand)] / um_count i32[1] receive the reference's unmatched_detections list
    in the reference's order; um_rows i32[>= len(row_cand)] / um_rows_count i32[1] (optional) its unmatched_tracks list.
    No host sync; returns the status tensor."""
    if not (isinstance(cost, torch.Tensor) and cost.is_cuda and cost.dtype == torch.float64 and cost.dim() == 2 and cost.stride(1) == 1):
        raise ValueError("lap_exact: a float64 CUDA matrix with unit column stride is required")
    dev = cost.device
    T, M = cost.shape
    for name, t in (("row_cand", row_cand), ("col_cand", col_cand), ("match_col_of_row", match_col_of_row),
                    ("det_matched", det_matched), ("um_dets", um_dets), ("um_count", um_count)):
        if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == torch.int32 and t.is_contiguous()):
            raise ValueError("lap_exact: %s must be a contiguous int32 CUDA tensor" % name)
    if match_col_of_row.numel() != T or det_matched.numel() != M or um_dets.numel() < col_cand.numel() or um_count.numel() < 1:
        raise ValueError("lap_exact: match_col_of_row i32[T], det_matched i32[M], um_dets i32[>= len(col_cand)], um_count i32[1]")
    if status is None:
        status = torch.zeros(1, dtype=torch.int32, device=dev)
    if T and M:
        with torch.cuda.device(dev):
            L.check(L.load().reidb200_lap_exact(cost.data_ptr(), T, M, cost.stride(0), row_cand.data_ptr(), row_cand.numel(),
                                                col_cand.data_ptr(), L.ptr(col_count), col_cand.numel(), float(max_distance),
                                                match_col_of_row.data_ptr(), det_matched.data_ptr(), um_dets.data_ptr(),
                                                um_count.data_ptr(), L.ptr(um_rows), L.ptr(um_rows_count), status.data_ptr(),
                                                L.stream_ptr()))
    return status


def lap_gated(cost, row_cand, max_distance, match_col_of_row, det_matched, status=None):
    """One min_cost_matching solve on the device cost matrix f64[T, M]: rows `row_cand` (i32, device) that are still
    unmatched x columns still unmatched; matches are recorded in match_col_of_row i32[T] / det_matched i32[M] in place.
    Returns the status tensor (i32[1], non-zero if a solve met non-finite costs); no host sync."""
    if not (isinstance(cost, torch.Tensor) and cost.is_cuda and cost.dtype == torch.float64 and cost.dim() == 2 and cost.stride(1) == 1):
        raise ValueError("lap_gated: a float64 CUDA matrix with unit column stride is required")
    dev = cost.device
    T, M = cost.shape
    row_cand = _i32(row_cand, dev)
    if status is None:
        status = torch.zeros(1, dtype=torch.int32, device=dev)
    if match_col_of_row.numel() != T or det_matched.numel() != M or match_col_of_row.dtype != torch.int32 or det_matched.dtype != torch.int32:
        raise ValueError("lap_gated: match_col_of_row i32[T] and det_matched i32[M] are required")
    if T and M and row_cand.numel():
        lib = L.load()
        ws = L.workspace(lib.reidb200_lap_workspace_bytes(T, M), dev)
        with torch.cuda.device(dev):
            L.check(lib.reidb200_lap_gated(cost.data_ptr(), T, M, cost.stride(0), row_cand.data_ptr(), row_cand.numel(),
                                           float(max_distance), match_col_of_row.data_ptr(), det_matched.data_ptr(), status.data_ptr(),
                                           ws.data_ptr(), ws.numel(), L.stream_ptr()))
    return status
