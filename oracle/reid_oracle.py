"""CPU oracle for the ReID matching hot path  --  TEST INFRASTRUCTURE ONLY.

This module is a plain numpy restatement of the reference algorithms
(SuperbTUM/real-time-ReID-tracking @ bb113fd).  It is the *checker* for the
CUDA path: only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import it.  The product
package (``real-time-reid-tracking_b200``) never does; it fails loudly when
the CUDA library is missing.

Parity status: PINNED.  Every function below is checked by
``oracle/make_golden.py`` against the reference's own Python code imported in
place from ``/root/reference`` on seeded inputs; the resulting vectors are
committed under ``tests/golden/`` and re-checked by ``tests/test_oracle.py``.
Two boundaries stay "parity unpinned" because the reference holds nothing to
pin them with:
  * the faiss k-NN (faiss-gpu 1.7.4 is not vendored / not installable): restated
    as exact fp32 squared-L2 brute force, ascending (``knn_l2``);
  * the lambda-blend of ``re_ranking`` (absent from the reference).

Each function cites the reference file:line it follows.
"""
from __future__ import annotations

import math

import numpy as np

INFTY_COST = 1e5  # boxmot/trackers/strongsort/sort/linear_assignment.py:9
CHI2INV95_4 = 9.4877  # boxmot/utils/matching.py:12-22 (4 dof)


# --------------------------------------------------------------------------
# B1/B2  distance matrices  (reid/losses/utils.py:12-35)
# --------------------------------------------------------------------------
def euclidean_dist(x: np.ndarray, y: np.ndarray) -> np.ndarray:
    """sqrt(clamp(|x|^2 + |y|^2 - 2 x.y^T, 1e-12)) in fp32 (reid/losses/utils.py:21-35)."""
    x = np.asarray(x, np.float32)
    y = np.asarray(y, np.float32)
    xx = (x * x).sum(1, dtype=np.float32)[:, None]
    yy = (y * y).sum(1, dtype=np.float32)[None, :]
    d = (xx + yy).astype(np.float32)
    d = d + np.float32(-2.0) * (x @ y.T)
    return np.sqrt(np.maximum(d, np.float32(1e-12)), dtype=np.float32)


def cosine_dist(x: np.ndarray, y: np.ndarray) -> np.ndarray:
    """(1 - x.y^T / (|x||y|)) / 2 in fp32 (reid/losses/utils.py:12-18)."""
    x = np.asarray(x, np.float32)
    y = np.asarray(y, np.float32)
    up = x @ y.T
    nx = np.sqrt((x * x).sum(1, dtype=np.float32))[:, None]
    ny = np.sqrt((y * y).sum(1, dtype=np.float32))[None, :]
    return ((np.float32(1.0) - up / (nx * ny)) / np.float32(2.0)).astype(np.float32)


def sql2_dist(xq: np.ndarray, xb: np.ndarray) -> np.ndarray:
    """Squared L2, the quantity faiss METRIC_L2 ranks by (reid/faiss_utils.py:56-118)."""
    xq = np.asarray(xq, np.float32)
    xb = np.asarray(xb, np.float32)
    qq = (xq * xq).sum(1, dtype=np.float32)[:, None]
    bb = (xb * xb).sum(1, dtype=np.float32)[None, :]
    return (qq + bb - np.float32(2.0) * (xq @ xb.T)).astype(np.float32)


# --------------------------------------------------------------------------
# C1-C3  brute-force k-NN  (faiss stand-in; reid/faiss_utils.py:30-139)
# --------------------------------------------------------------------------
def knn_l2(xb: np.ndarray, xq: np.ndarray, k: int, block: int = 2048):
    """Exact squared-L2 top-k, ascending; ties broken by lower index.

    Stands in for faiss ``bfKnn`` / ``GpuIndexFlatL2`` / ``IndexFlatL2`` /
    ``IndexShards`` (reid/faiss_utils.py:97-111,121-139).  Returns (D f32[nq,k],
    I i64[nq,k]).  "parity unpinned" (faiss absent).
    """
    xb = np.asarray(xb, np.float32)
    xq = np.asarray(xq, np.float32)
    nq = xq.shape[0]
    D = np.empty((nq, k), np.float32)
    I = np.empty((nq, k), np.int64)
    for s in range(0, nq, block):
        d = sql2_dist(xq[s:s + block], xb)
        idx = np.argsort(d, axis=1, kind="stable")[:, :k]
        I[s:s + block] = idx
        D[s:s + block] = np.take_along_axis(d, idx, 1)
    return D, I


# --------------------------------------------------------------------------
# D1-D7  k-reciprocal Jaccard distance  (reid/faiss_utils.py:142-244)
# --------------------------------------------------------------------------
def k_reciprocal_neigh(initial_rank: np.ndarray, i: int, k1: int) -> np.ndarray:
    """R(i,k1): forward neighbours whose own forward list holds i (faiss_utils.py:142-146)."""
    fwd = initial_rank[i, :k1 + 1]
    bwd = initial_rank[fwd, :k1 + 1]
    keep = np.where(bwd == i)[0]
    return fwd[keep]


def _softmax_f32(v: np.ndarray) -> np.ndarray:
    """Row softmax in fp32 the way torch.softmax does it: exp(v - max) / sum."""
    v = v.astype(np.float32)
    e = np.exp(v - v.max(), dtype=np.float32)
    return (e / e.sum(dtype=np.float32)).astype(np.float32)


def jaccard_sparse_rows(feats: np.ndarray, initial_rank: np.ndarray, k1: int, k2: int):
    """Stages D3-D5 in sparse form: list of (cols i64 sorted, vals f32) per row after
    local expansion, softmax weighting and k2 query expansion (faiss_utils.py:183-214)."""
    feats = np.asarray(feats, np.float32)
    N = feats.shape[0]
    half = int(np.around(k1 / 2))
    r_full = [k_reciprocal_neigh(initial_rank, i, k1) for i in range(N)]
    r_half = [k_reciprocal_neigh(initial_rank, i, half) for i in range(N)]
    rows = []
    for i in range(N):
        base = r_full[i]
        grown = base
        for c in base:
            rc = r_half[c]
            if len(np.intersect1d(rc, base)) > 2 / 3 * len(rc):
                grown = np.append(grown, rc)
        cols = np.unique(grown)
        if cols.size:
            dist = np.float32(2.0) - np.float32(2.0) * (feats[i][None, :] @ feats[cols].T)[0]
            vals = _softmax_f32(-dist)
        else:
            vals = np.zeros((0,), np.float32)
        rows.append((cols.astype(np.int64), vals))
    if k2 != 1:
        out = []
        for i in range(N):
            acc: dict[int, np.float32] = {}
            for j in initial_rank[i, :k2]:
                cj, vj = rows[j]
                for c, v in zip(cj.tolist(), vj):
                    acc[c] = np.float32(acc.get(c, np.float32(0.0)) + v)
            cols = np.array(sorted(acc), np.int64)
            vals = np.array([acc[c] for c in cols.tolist()], np.float32) / np.float32(len(initial_rank[i, :k2]))
            out.append((cols, vals.astype(np.float32)))
        rows = out
    return rows


def compute_jaccard_distance(feats: np.ndarray, k1: int = 20, k2: int = 6, use_float16: bool = False,
                             initial_rank: np.ndarray | None = None) -> np.ndarray:
    """Dense N x N Jaccard distance, same algorithm and dense temporaries as the
    reference (reid/faiss_utils.py:149-244) so that timing it is a fair CPU baseline.

    ``initial_rank`` (N,k1) may be supplied (any faiss search_option gives the same
    thing); otherwise the exact k-NN stand-in is used with k = k1 (faiss_utils.py:162).
    """
    feats = np.asarray(feats, np.float32)
    N = feats.shape[0]
    mat = np.float16 if use_float16 else np.float32
    if initial_rank is None:
        _, initial_rank = knn_l2(feats, feats, k1)
    half = int(np.around(k1 / 2))
    r_full, r_half = [], []
    for i in range(N):
        r_full.append(k_reciprocal_neigh(initial_rank, i, k1))
        r_half.append(k_reciprocal_neigh(initial_rank, i, half))
    V = np.zeros((N, N), mat)
    for i in range(N):
        base = r_full[i]
        grown = base
        for c in base:
            rc = r_half[c]
            if len(np.intersect1d(rc, base)) > 2 / 3 * len(rc):
                grown = np.append(grown, rc)
        cols = np.unique(grown)
        dist = np.float32(2.0) - np.float32(2.0) * (feats[i][None, :] @ feats[cols].T)[0]
        if cols.size:
            V[i, cols] = _softmax_f32(-dist).astype(mat)
    if k2 != 1:
        Vq = np.zeros_like(V)
        for i in range(N):
            Vq[i] = np.mean(V[initial_rank[i, :k2]], axis=0)
        V = Vq
    inv = [np.where(V[:, j] != 0)[0] for j in range(N)]
    J = np.zeros((N, N), mat)
    for i in range(N):
        acc = np.zeros((N,), mat)
        nz = np.where(V[i] != 0)[0]
        for j in nz:
            rows = inv[j]
            acc[rows] = acc[rows] + np.minimum(V[i, j], V[rows, j])
        J[i] = 1 - acc / (2 - acc)
    J[J < 0] = 0.0
    return J


def re_ranking_from_feats(qf: np.ndarray, gf: np.ndarray, k1: int = 20, k2: int = 6,
                          lambda_value: float = 0.3) -> np.ndarray:
    """north_star addition ("parity unpinned" for the blend):
    merged = cat(gallery, query) as reid/image_reid_inference.py:270 does, Jaccard
    on the merged set (D2-D7), then ``(1-lambda)*J[q,g] + lambda*euclidean_dist(q,g)``."""
    G = gf.shape[0]
    merged = np.concatenate([gf, qf], 0).astype(np.float32)
    J = compute_jaccard_distance(merged, k1, k2).astype(np.float32)
    orig = euclidean_dist(qf, gf)
    lam = np.float32(lambda_value)
    return (J[G:, :G] * (np.float32(1.0) - lam) + orig * lam).astype(np.float32)


# --------------------------------------------------------------------------
# B3-B6  CMC / mAP  (reid/evaluate.py:33-105)
# --------------------------------------------------------------------------
def compute_mAP(index: np.ndarray, good_index: np.ndarray, junk_index: np.ndarray):
    """AP (python float) and cmc int32[len(index)], sentinel cmc[0] = -1 (evaluate.py:78-105)."""
    ap = 0
    cmc = np.zeros(len(index), np.int32)
    if good_index.size == 0:
        cmc[0] = -1
        return ap, cmc
    index = index[np.isin(index, junk_index, invert=True)]
    ngood = len(good_index)
    rows_good = np.argwhere(np.isin(index, good_index)).flatten()
    cmc[rows_good[0]:] = 1
    for i in range(ngood):
        d_recall = 1.0 / ngood
        precision = (i + 1) * 1.0 / (rows_good[i] + 1)
        old_precision = i * 1.0 / rows_good[i] if rows_good[i] != 0 else 1.0
        ap = ap + d_recall * (old_precision + precision) / 2
    return ap, cmc


def _good_junk(ql, qc, gl, gc):
    same_id = np.argwhere(gl == ql)
    same_cam = np.argwhere(gc == qc)
    good = np.setdiff1d(same_id, same_cam, assume_unique=True)
    junk = np.append(np.intersect1d(same_id, same_cam), np.argwhere(gl == -1))
    return good, junk


def evaluate(qf, ql, qc, gf, gl, gc):
    """One query: score = gf.q, descending (evaluate.py:55-75)."""
    score = (np.asarray(gf, np.float32) @ np.asarray(qf, np.float32).reshape(-1, 1))[:, 0]
    index = np.argsort(score)[::-1]
    good, junk = _good_junk(ql, qc, np.asarray(gl), np.asarray(gc))
    return compute_mAP(index, good, junk)


def evaluate_all(qf, ql, qc, gf, gl, gc):
    """CMC float32[G] and mAP float over all queries, divided by Q (evaluate.py:33-52)."""
    Q = qf.shape[0]
    CMC = np.zeros(gf.shape[0], np.int32)
    ap = 0.0
    for i in range(Q):
        ap_i, cmc_i = evaluate(qf[i], ql[i], qc[i], gf, gl, gc)
        if cmc_i[0] == -1:
            continue
        CMC = CMC + cmc_i
        ap += ap_i
    return CMC.astype(np.float32) / np.float32(Q), ap / Q


def evaluate_all_from_dist(dist, ql, qc, gl, gc):
    """Same reduction on a precomputed Q x G *distance* matrix (ascending = better);
    used for the re-ranked evaluation (image_reid_inference.py:284-315 ordering)."""
    Q, G = dist.shape
    CMC = np.zeros(G, np.int32)
    ap = 0.0
    gl = np.asarray(gl)
    gc = np.asarray(gc)
    for i in range(Q):
        index = np.argsort(dist[i], kind="stable")
        good, junk = _good_junk(ql[i], qc[i], gl, gc)
        ap_i, cmc_i = compute_mAP(index, good, junk)
        if cmc_i[0] == -1:
            continue
        CMC = CMC + cmc_i
        ap += ap_i
    return CMC.astype(np.float32) / np.float32(Q), ap / Q


def evaluate_per_query_from_dist(dist, ql, qc, gl, gc):
    """Per-query (AP f64[Q], first-hit rank i32[Q] with -1 = the cmc[0] == -1 sentinel) of a DISTANCE matrix: what
    reid/evaluate.py:55-75 evaluate() returns for each query, on the ranking evaluate_all_from_dist uses."""
    Q = dist.shape[0]
    gl = np.asarray(gl)
    gc = np.asarray(gc)
    aps = np.zeros(Q)
    first = np.full(Q, -1, np.int32)
    for i in range(Q):
        index = np.argsort(dist[i], kind="stable")
        good, junk = _good_junk(ql[i], qc[i], gl, gc)
        ap_i, cmc_i = compute_mAP(index, good, junk)
        if cmc_i[0] == -1:
            continue
        aps[i] = ap_i
        first[i] = int(np.argmax(cmc_i))
    return aps, first


# --------------------------------------------------------------------------
# A1/A2  crop + cv2-style bilinear resize + normalise
# (modification_deepsort/feature_extractor.py:24-46; strong_sort.py:115-121,149-159)
# --------------------------------------------------------------------------
def xywh_to_xyxy(box, width: int, height: int):
    """Centre-xywh -> int-truncated, clamped xyxy (strong_sort.py:115-121)."""
    x, y, w, h = [float(v) for v in box]
    x1 = max(int(x - w / 2), 0)
    x2 = min(int(x + w / 2), width - 1)
    y1 = max(int(y - h / 2), 0)
    y2 = min(int(y + h / 2), height - 1)
    return x1, y1, x2, y2


def _cv_linear_taps(src: int, dst: int):
    """Source index pair + fp32 weights of cv2.resize INTER_LINEAR (half-pixel centres,
    no antialias).  ``clamp_w`` is True on the horizontal axis where OpenCV zeroes the
    fraction at the borders; on the vertical axis it only clamps the row index."""
    scale = 1.0 / (dst / src)
    d = np.arange(dst)
    f = ((d + 0.5) * scale - 0.5).astype(np.float32)
    s = np.floor(f).astype(np.int64)
    f = (f - s.astype(np.float32)).astype(np.float32)
    return s, f


def resize_linear_cv(img: np.ndarray, out_w: int, out_h: int) -> np.ndarray:
    """cv2.resize(img_f32, (out_w,out_h)) for HxWxC float32 (INTER_LINEAR)."""
    img = np.asarray(img, np.float32)
    h, w = img.shape[:2]
    sx, fx = _cv_linear_taps(w, out_w)
    sy, fy = _cv_linear_taps(h, out_h)
    # horizontal: fraction forced to 0 outside [0, w-1)
    lo = sx < 0
    hi = sx >= w - 1
    fx = np.where(lo | hi, np.float32(0.0), fx).astype(np.float32)
    sx0 = np.where(lo, 0, np.where(hi, w - 1, sx))
    sx1 = np.minimum(sx0 + 1, w - 1)
    a0 = (np.float32(1.0) - fx)[None, :, None]
    a1 = fx[None, :, None]
    rows = img[:, sx0, :] * a0 + img[:, sx1, :] * a1  # [h, out_w, C]
    # vertical: indices clamped, weights kept
    sy0 = np.clip(sy, 0, h - 1)
    sy1 = np.clip(sy + 1, 0, h - 1)
    b0 = (np.float32(1.0) - fy)[:, None, None]
    b1 = fy[:, None, None]
    return (rows[sy0] * b0 + rows[sy1] * b1).astype(np.float32)


def preprocess_deepsort(crops, out_w: int = 128, out_h: int = 256, mean=(0.5, 0.5, 0.5), std=(0.5, 0.5, 0.5)):
    """Extractor._preprocess: u8 HWC crops -> f32 [M,3,out_h,out_w] (feature_extractor.py:31-46)."""
    out = np.empty((len(crops), 3, out_h, out_w), np.float32)
    m = np.asarray(mean, np.float32)[:, None, None]
    s = np.asarray(std, np.float32)[:, None, None]
    for i, im in enumerate(crops):
        r = resize_linear_cv(im.astype(np.float32) / np.float32(255.0), out_w, out_h)
        out[i] = (r.transpose(2, 0, 1) - m) / s
    return out


# --------------------------------------------------------------------------
# A3  PIL antialiased bilinear (boxmot/appearance/reid_multibackend.py:58-66,210-220)
# third-party algorithm: Pillow 12.2 ImagingResample (8bpc fixed point, two passes)
# --------------------------------------------------------------------------
_PIL_PREC = 32 - 8 - 2


def pil_bilinear_coeffs(in_size: int, out_size: int):
    """Per-output bounds (xmin, n) and int32 fixed-point taps, Pillow precompute_coeffs +
    normalize_coeffs_8bpc with the bilinear (triangle) filter, support scaled when shrinking."""
    scale = in_size / out_size
    fscale = max(scale, 1.0)
    support = 1.0 * fscale
    ksize = int(math.ceil(support)) * 2 + 1
    bounds = np.zeros((out_size, 2), np.int32)
    kk = np.zeros((out_size, ksize), np.int32)
    inv = 1.0 / fscale
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        xmin = int(center - support + 0.5)
        xmin = max(xmin, 0)
        xmax = int(center + support + 0.5)
        xmax = min(xmax, in_size)
        n = xmax - xmin
        w = np.empty(n, np.float64)
        for x in range(n):
            a = abs((x + xmin - center + 0.5) * inv)
            w[x] = 1.0 - a if a < 1.0 else 0.0
        ww = w.sum() if n else 0.0
        # Pillow accumulates ww sequentially; keep the same order
        ww = 0.0
        for x in range(n):
            ww += w[x]
        if ww != 0.0:
            w = w / ww
        for x in range(n):
            v = w[x] * (1 << _PIL_PREC)
            kk[xx, x] = int(v - 0.5) if w[x] < 0 else int(v + 0.5)
        bounds[xx] = (xmin, n)
    return bounds, kk


def _pil_clip8(v: np.ndarray) -> np.ndarray:
    return np.clip(v >> _PIL_PREC, 0, 255).astype(np.uint8)


def resize_pil_bilinear_u8(img: np.ndarray, out_h: int, out_w: int) -> np.ndarray:
    """PIL.Image.resize((out_w,out_h), BILINEAR) on an HxWx3 uint8 image: horizontal pass
    then vertical pass, uint8 intermediate."""
    img = np.asarray(img, np.uint8)
    h, w = img.shape[:2]
    cur = img
    if out_w != w:
        bounds, kk = pil_bilinear_coeffs(w, out_w)
        tmp = np.empty((h, out_w, 3), np.uint8)
        for xx in range(out_w):
            x0, n = bounds[xx]
            acc = (cur[:, x0:x0 + n, :].astype(np.int64) * kk[xx, :n][None, :, None]).sum(1) + (1 << (_PIL_PREC - 1))
            tmp[:, xx, :] = _pil_clip8(acc)
        cur = tmp
    if out_h != h:
        bounds, kk = pil_bilinear_coeffs(h, out_h)
        tmp = np.empty((out_h, cur.shape[1], 3), np.uint8)
        for yy in range(out_h):
            y0, n = bounds[yy]
            acc = (cur[y0:y0 + n].astype(np.int64) * kk[yy, :n][:, None, None]).sum(0) + (1 << (_PIL_PREC - 1))
            tmp[yy] = _pil_clip8(acc)
        cur = tmp
    return cur


def preprocess_boxmot(crops, out_h: int = 256, out_w: int = 128,
                      mean=(0.485, 0.456, 0.406), std=(0.229, 0.224, 0.225)):
    """ReIDDetectMultiBackend._preprocess: ToPILImage -> Resize((256,128)) -> ToTensor ->
    Normalize(ImageNet) (reid_multibackend.py:58-66,210-220)."""
    out = np.empty((len(crops), 3, out_h, out_w), np.float32)
    m = np.asarray(mean, np.float32)[:, None, None]
    s = np.asarray(std, np.float32)[:, None, None]
    for i, im in enumerate(crops):
        r = resize_pil_bilinear_u8(im, out_h, out_w).astype(np.float32) / np.float32(255.0)
        out[i] = (r.transpose(2, 0, 1) - m) / s
    return out


# --------------------------------------------------------------------------
# E2-E4  nearest-neighbour appearance metric  (boxmot/utils/matching.py:255-409)
# --------------------------------------------------------------------------
def pdist(a, b):
    """Squared L2, clipped at 0 (matching.py:255-275)."""
    a, b = np.asarray(a), np.asarray(b)
    if len(a) == 0 or len(b) == 0:
        return np.zeros((len(a), len(b)))
    a2, b2 = np.square(a).sum(axis=1), np.square(b).sum(axis=1)
    return np.clip(-2.0 * np.dot(a, b.T) + a2[:, None] + b2[None, :], 0.0, float(np.inf))


def cosine_distance(a, b, data_is_normalized: bool = False):
    """1 - a^.b^T with both sides re-normalised in their own dtype (matching.py:278-298)."""
    a = np.asarray(a)
    b = np.asarray(b)
    if not data_is_normalized:
        a = a / np.linalg.norm(a, axis=1, keepdims=True)
        b = b / np.linalg.norm(b, axis=1, keepdims=True)
    return 1.0 - np.dot(a, b.T)


def nn_cosine_distance(samples, feats):
    """min over a track's samples of the cosine distance (matching.py:321-339)."""
    return cosine_distance(np.asarray(samples), np.asarray(feats)).min(axis=0)


def nn_distance_matrix(samples_per_track, feats):
    """NearestNeighborDistanceMetric.distance: f64 [T,M] holding fp32 values (matching.py:391-409)."""
    cost = np.zeros((len(samples_per_track), len(feats)))
    for t, s in enumerate(samples_per_track):
        cost[t, :] = nn_cosine_distance(s, feats)
    return cost


# --------------------------------------------------------------------------
# E6/E7/E13  Mahalanobis gate, blend, clamp
# --------------------------------------------------------------------------
def tlwh_to_xyah_f32(tlwh):
    """Detection.to_xyah in float32 (sort/detection.py:29-50)."""
    r = np.asarray(tlwh, np.float32).copy()
    r[..., :2] += r[..., 2:] / np.float32(2)
    r[..., 2] /= r[..., 3]
    return r


def gating_distance(x4, SI, measurements):
    """sqrt(y^T SI y), y = z - x[:4], with the track's *stale* SI
    (strongsort_kf_adapter.py:137-169 -> kalman_filter.py:577-586)."""
    x4 = np.asarray(x4, np.float64).reshape(4, 1)
    SI = np.asarray(SI, np.float64).reshape(4, 4)
    out = np.zeros((len(measurements),))
    for i, z in enumerate(measurements):
        y = np.asarray(z).reshape(4, 1) - x4
        out[i] = math.sqrt(float(np.dot(np.dot(y.T, SI), y)[0, 0]))
    return out


def gate_cost_matrix(cost, x4s, SIs, measurements, mc_lambda, gated_cost=INFTY_COST):
    """In place: cost[row, gd > 9.4877] = 1e5; cost[row] = l*cost[row] + (1-l)*gd
    (sort/linear_assignment.py:147-198)."""
    for row in range(cost.shape[0]):
        gd = gating_distance(x4s[row], SIs[row], measurements)
        cost[row, gd > CHI2INV95_4] = gated_cost
        cost[row] = mc_lambda * cost[row] + (1 - mc_lambda) * gd
    return cost


def clamp_cost(cost, max_distance):
    """cost[cost > max_distance] = max_distance + 1e-5 (sort/linear_assignment.py:58-59)."""
    cost[cost > max_distance] = max_distance + 1e-5
    return cost


# --------------------------------------------------------------------------
# E9-E12  IoU / DIoU cost
# --------------------------------------------------------------------------
def track_to_tlwh(mean4):
    """Track.to_tlwh from the KF mean (xyah) in fp64 (sort/track.py:107-120)."""
    r = np.asarray(mean4, np.float64).copy()
    r[2] *= r[3]
    r[:2] -= r[2:] / 2
    return r


def iou(bbox, candidates):
    """Plain IoU of one tlwh box vs M tlwh boxes (strongsort/sort/iou_matching.py:9-45)."""
    b_tl, b_br = bbox[:2], bbox[:2] + bbox[2:]
    c_tl = candidates[:, :2]
    c_br = candidates[:, :2] + candidates[:, 2:]
    tl = np.stack([np.maximum(b_tl[0], c_tl[:, 0]), np.maximum(b_tl[1], c_tl[:, 1])], 1)
    br = np.stack([np.minimum(b_br[0], c_br[:, 0]), np.minimum(b_br[1], c_br[:, 1])], 1)
    wh = np.maximum(0.0, br - tl)
    inter = wh.prod(axis=1)
    return inter / (bbox[2:].prod() + candidates[:, 2:].prod(axis=1) - inter)


def diou(bbox, candidates):
    """IoU - centre^2 / outer-diagonal^2 (modification_deepsort/iou_matching.py:5-47)."""
    b_tl, b_br = bbox[:2], bbox[:2] + bbox[2:]
    c_tl = candidates[:, :2]
    c_br = candidates[:, :2] + candidates[:, 2:]
    b_c = np.asarray([(b_tl[1] + b_br[1]) / 2, (b_tl[0] + b_br[0]) / 2])
    c_c = np.asarray([(c_tl[:, 1] + c_br[:, 1]) / 2, (c_tl[:, 0] + c_br[:, 0]) / 2]).T
    d = np.sum((b_c - c_c) ** 2, axis=1)
    o_tl = np.stack([np.minimum(b_tl[0], c_tl[:, 0]), np.minimum(b_tl[1], c_tl[:, 1])], 1)
    o_br = np.stack([np.maximum(b_br[0], c_br[:, 0]), np.maximum(b_br[1], c_br[:, 1])], 1)
    rou = np.sum((o_tl - o_br) ** 2, axis=1)
    return iou(bbox, candidates) - d / rou


def iou_cost(track_tlwh, time_since_update, det_tlwh, use_diou: bool = False):
    """f64 [T,M]: rows with time_since_update > 1 = 1e5 else 1 - iou
    (strongsort/sort/iou_matching.py:48-86)."""
    T, M = len(track_tlwh), len(det_tlwh)
    cost = np.zeros((T, M))
    fn = diou if use_diou else iou
    cands = np.asarray(det_tlwh)
    for r in range(T):
        if time_since_update[r] > 1:
            cost[r, :] = INFTY_COST
            continue
        cost[r, :] = 1.0 - fn(np.asarray(track_tlwh[r]), cands)
    return cost


# --------------------------------------------------------------------------
# E13 + N1 boundary: Hungarian on the clamped cost, and the two-stage _match
# --------------------------------------------------------------------------
def min_cost_matching_from_cost(cost, max_distance, track_indices, detection_indices):
    """sort/linear_assignment.py:12-77 given the metric's cost matrix (rows = track_indices,
    cols = detection_indices).  Clamp, scipy Hungarian, threshold filter."""
    from scipy.optimize import linear_sum_assignment

    track_indices = list(track_indices)
    detection_indices = list(detection_indices)
    if len(detection_indices) == 0 or len(track_indices) == 0:
        return [], track_indices, detection_indices
    cost = np.array(cost, np.float64, copy=True)
    cost[cost > max_distance] = max_distance + 1e-5
    rows, cols = linear_sum_assignment(cost)
    matches, um_t, um_d = [], [], []
    for c, d in enumerate(detection_indices):
        if c not in cols:
            um_d.append(d)
    for r, t in enumerate(track_indices):
        if r not in rows:
            um_t.append(t)
    for r, c in zip(rows, cols):
        t, d = track_indices[r], detection_indices[c]
        if cost[r, c] > max_distance:
            um_t.append(t)
            um_d.append(d)
        else:
            matches.append((t, d))
    return matches, um_t, um_d


def match_frame(fr, max_dist, max_iou_dist, mc_lambda):
    """Tracker._match (sort/tracker.py:175-233) on one recorded frame (see
    oracle/make_golden.py::_mk_tracker for the record layout)."""
    T, M = len(fr["trk_id"]), len(fr["det_tlwh"])
    ids = fr["sample_ids"].tolist()
    off = np.concatenate([[0], np.cumsum(fr["sample_cnt"])])
    samples = {k: fr["sample_rows"][off[i]:off[i + 1]] for i, k in enumerate(ids)}
    confirmed = [i for i in range(T) if fr["trk_confirmed"][i]]
    unconfirmed = [i for i in range(T) if not fr["trk_confirmed"][i]]
    dets = list(range(M))
    if confirmed and M:
        cost = nn_distance_matrix([samples[int(fr["trk_id"][i])] for i in confirmed], fr["det_feat"])
        cost = gate_cost_matrix(cost, fr["trk_mean"][confirmed, :4], fr["trk_SI"][confirmed],
                                tlwh_to_xyah_f32(fr["det_tlwh"]), mc_lambda)
    else:
        cost = np.zeros((len(confirmed), M))
    m_a, um_a, um_d = min_cost_matching_from_cost(cost, max_dist, confirmed, dets)
    um_a = list(set(confirmed) - set(k for k, _ in m_a))  # matching_cascade :143
    cand = unconfirmed + [k for k in um_a if fr["trk_tsu"][k] == 1]
    um_a = [k for k in um_a if fr["trk_tsu"][k] != 1]
    if cand and um_d:
        tl = [track_to_tlwh(fr["trk_mean"][k, :4]) for k in cand]
        ic = iou_cost(tl, fr["trk_tsu"][cand], fr["det_tlwh"][um_d])
    else:
        ic = np.zeros((len(cand), len(um_d)))
    m_b, um_b, um_d = min_cost_matching_from_cost(ic, max_iou_dist, cand, um_d)
    return m_a + m_b, list(set(um_a + um_b)), um_d


# --------------------------------------------------------------------------
# N2  the other boxmot cost families: pairwise box overlaps of
# boxmot/utils/iou.py:4-164 (xyxy boxes, arithmetic in the inputs' dtype), the
# "+1" pixel IoU of boxmot/utils/matching.py:216-248 (float32 inputs), and the
# score / IoU fusions of matching.py:189-213.
# --------------------------------------------------------------------------
def _pair_terms(b1, b2):
    """Broadcast views (N,1) x (1,K) of the corners plus the overlap area and plain IoU."""
    a = np.asarray(b1)[:, None, :]
    b = np.asarray(b2)[None, :, :]
    iw = np.maximum(0.0, np.minimum(a[..., 2], b[..., 2]) - np.maximum(a[..., 0], b[..., 0]))
    ih = np.maximum(0.0, np.minimum(a[..., 3], b[..., 3]) - np.maximum(a[..., 1], b[..., 1]))
    inter = iw * ih
    union = (a[..., 2] - a[..., 0]) * (a[..., 3] - a[..., 1]) + (b[..., 2] - b[..., 0]) * (b[..., 3] - b[..., 1]) - inter
    return a, b, inter, inter / union


def box_pairs(b1, b2, mode: str = "iou"):
    """iou_batch / giou_batch / diou_batch / ciou_batch (boxmot/utils/iou.py:4-23, 26-59, 62-102, 105-164):
    [N,K] matrix; giou/diou/ciou are rescaled to (0,1) by (x + 1) / 2 like the reference."""
    a, b, inter, iou = _pair_terms(b1, b2)
    if mode == "iou":
        return iou
    ex1, ey1 = np.minimum(a[..., 0], b[..., 0]), np.minimum(a[..., 1], b[..., 1])
    ex2, ey2 = np.maximum(a[..., 2], b[..., 2]), np.maximum(a[..., 3], b[..., 3])
    if mode == "giou":
        enclose = (ex2 - ex1) * (ey2 - ey1)
        return (iou - (enclose - inter) / enclose + 1.0) / 2.0
    cx1, cy1 = (a[..., 0] + a[..., 2]) / 2.0, (a[..., 1] + a[..., 3]) / 2.0
    cx2, cy2 = (b[..., 0] + b[..., 2]) / 2.0, (b[..., 1] + b[..., 3]) / 2.0
    inner = (cx1 - cx2) ** 2 + (cy1 - cy2) ** 2
    outer = (ex2 - ex1) ** 2 + (ey2 - ey1) ** 2
    if mode == "diou":
        return (iou - inner / outer + 1) / 2.0
    if mode == "ciou":
        w1, h1 = a[..., 2] - a[..., 0], a[..., 3] - a[..., 1] + 1.0
        w2, h2 = b[..., 2] - b[..., 0], b[..., 3] - b[..., 1] + 1.0
        at = np.arctan(w2 / h2) - np.arctan(w1 / h1)
        v = (4 / (np.pi ** 2)) * (at ** 2)
        alpha = v / ((1 - iou) + v)
        return (iou - inner / outer - alpha * v + 1) / 2.0
    raise ValueError(mode)


def bbox_ious(boxes, query_boxes):
    """The '+1' pixel-inclusive IoU of boxmot/utils/matching.py:216-248 on float32 tlbr boxes: widths/heights and the
    overlap are float32, the union is promoted to float64 by `float(...)`, the quotient is stored as float32."""
    bx = np.asarray(boxes, np.float32)
    qb = np.asarray(query_boxes, np.float32)
    one = np.float32(1)
    a = bx[:, None, :]
    q = qb[None, :, :]
    iw = np.minimum(a[..., 2], q[..., 2]) - np.maximum(a[..., 0], q[..., 0]) + one
    ih = np.minimum(a[..., 3], q[..., 3]) - np.maximum(a[..., 1], q[..., 1]) + one
    qa = (q[..., 2] - q[..., 0] + one) * (q[..., 3] - q[..., 1] + one)
    ba = (a[..., 2] - a[..., 0] + one) * (a[..., 3] - a[..., 1] + one)
    inter = iw * ih                                   # float32
    ua = ((ba + qa) - inter).astype(np.float64)       # float32 arithmetic, then float(...)
    out = np.zeros(inter.shape, np.float32)
    ok = (iw > 0) & (ih > 0)
    out[ok] = (inter.astype(np.float64)[ok] / ua[ok]).astype(np.float32)
    return out


def iou_distance(atlbrs, btlbrs):
    """1 - bbox_ious (matching.py:71-114), float32; empty inputs give an empty float32 matrix."""
    if len(atlbrs) == 0 or len(btlbrs) == 0:
        return np.zeros((len(atlbrs), len(btlbrs)), np.float32)
    return 1 - bbox_ious(atlbrs, btlbrs)


def fuse_score(cost, det_scores):
    """matching.py:203-213: 1 - (1 - cost) * score, float64 through the float64 score vector."""
    if np.asarray(cost).size == 0:
        return cost
    return 1 - (1 - np.asarray(cost)) * np.asarray(det_scores, np.float64)[None, :]


def fuse_iou(cost, atlbrs, btlbrs):
    """matching.py:189-200: 1 - (1 - cost) * (1 + iou_sim) / 2 (the score factor is commented out in the reference)."""
    if np.asarray(cost).size == 0:
        return cost
    iou_sim = 1 - iou_distance(atlbrs, btlbrs)
    return 1 - (1 - np.asarray(cost)) * (1 + iou_sim) / 2


def embedding_distance(track_feats, det_feats):
    """matching.py:138-160 with metric='cosine': max(0, 1 - u.v / (|u| |v|)) in float64 on float32 inputs
    (scipy.spatial.distance.cdist promotes to double)."""
    t = np.asarray(track_feats, np.float32).astype(np.float64)
    d = np.asarray(det_feats, np.float32).astype(np.float64)
    if t.size == 0 or d.size == 0:
        return np.zeros((len(t), len(d)), np.float32)
    sim = (t @ d.T) / (np.linalg.norm(t, axis=1)[:, None] * np.linalg.norm(d, axis=1)[None, :])
    return np.maximum(0.0, 1.0 - sim)


def ocsort_cost(detections, trackers, velocities, previous_obs, vdc_weight):
    """(iou_matrix [D,T], angle_diff_cost [D,T]) of OC-SORT's associate (boxmot/utils/association.py:6-16 and
    :122-140), float64: the two terms of final_cost = -(iou + angle [+ emb])."""
    det = np.asarray(detections, np.float64)
    trk = np.asarray(trackers, np.float64)
    vel = np.asarray(velocities, np.float64)
    prev = np.asarray(previous_obs, np.float64)
    cx1, cy1 = (det[:, 0] + det[:, 2]) / 2.0, (det[:, 1] + det[:, 3]) / 2.0          # [D]
    cx2, cy2 = (prev[:, 0] + prev[:, 2])[:, None] / 2.0, (prev[:, 1] + prev[:, 3])[:, None] / 2.0  # [T,1]
    dx, dy = cx1 - cx2, cy1 - cy2                                                       # [T,D]
    norm = np.sqrt(dx ** 2 + dy ** 2) + 1e-6
    dx, dy = dx / norm, dy / norm
    cosang = np.clip(vel[:, 1:2] * dx + vel[:, 0:1] * dy, -1, 1)
    diff = (np.pi / 2.0 - np.abs(np.arccos(cosang))) / np.pi
    valid = np.where(prev[:, 4] < 0, 0.0, 1.0)[:, None]
    angle = (((valid * diff) * vdc_weight).T) * det[:, -1][:, None]
    return box_pairs(det[:, :4], trk[:, :4], "iou"), angle


# --------------------------------------------------------------------------
# N3  StrongSORT's Kalman filter (boxmot/motion/adapters/strongsort_kf_adapter.py:25-135 on
# boxmot/motion/kalman_filter.py:340-379, 486-575), one track at a time, numpy float64.
# --------------------------------------------------------------------------
KF_W_POS, KF_W_VEL = 1.0 / 20, 1.0 / 160


def kf_initiate(measurement):
    m = np.asarray(measurement, np.float64)
    std = [2 * KF_W_POS * m[0], 2 * KF_W_POS * m[1], 1 * m[2], 2 * KF_W_POS * m[3],
           10 * KF_W_VEL * m[0], 10 * KF_W_VEL * m[1], 0.1 * m[2], 10 * KF_W_VEL * m[3]]
    return np.r_[m, np.zeros(4)], np.diag(np.square(std)), np.zeros((4, 4))


def kf_predict(x, P, dt=1.0):
    F = np.eye(8)
    for i in range(4):
        F[i, 4 + i] = dt
    std = [KF_W_POS * x[0], KF_W_POS * x[1], 1 * x[2], KF_W_POS * x[3], KF_W_VEL * x[0], KF_W_VEL * x[1], 0.1 * x[2], KF_W_VEL * x[3]]
    Q = np.diag(np.square(std))
    return F @ x, (F @ P) @ F.T + Q


def kf_update(x, P, z, confidence=0.0):
    """Returns (x, P, SI): SI is the inverse innovation covariance the NEXT frames gate with."""
    H = np.eye(4, 8)
    std = [(1 - confidence) * s for s in (KF_W_POS * x[3], KF_W_POS * x[3], 1e-1, KF_W_POS * x[3])]
    R = np.diag(np.square(std))
    y = np.asarray(z, np.float64) - H @ x
    PHT = P @ H.T
    SI = np.linalg.inv(H @ PHT + R)
    K = PHT @ SI
    A = np.eye(8) - K @ H
    return x + K @ y, (A @ P) @ A.T + (K @ R) @ K.T, SI


# ------------------------------------------------------------------------------------------
# N4: the steps around compute_jaccard_distance in reid/image_reid_inference.py
# ------------------------------------------------------------------------------------------
def tta_merge_normalize(e1: np.ndarray, e2: np.ndarray) -> np.ndarray:
    """(e1 + e2) / 2.0 then F.normalize(dim=1) = x / max(||x||_2, 1e-12) (image_reid_inference.py:252-253, 268-269)."""
    v = (np.asarray(e1, np.float32) + np.asarray(e2, np.float32)) / np.float32(2.0)
    nrm = np.sqrt((v * v).sum(1, dtype=np.float32, keepdims=True)).astype(np.float32)
    return (v / np.maximum(nrm, np.float32(1e-12))).astype(np.float32)


def clamp_add(dists: np.ndarray, other: np.ndarray = None) -> np.ndarray:
    """dists[dists < 0] = 0 ; dists += attribute_dist (image_reid_inference.py:287-290), float32."""
    d = np.array(dists, dtype=np.float32, copy=True)
    d[d < 0] = 0.0
    if other is not None:
        d += np.asarray(other, np.float32)
    return d


def dbscan_precomputed(dists: np.ndarray, eps: float, min_samples: int):
    """sklearn.cluster.DBSCAN(eps, min_samples, metric="precomputed").fit_predict(dists) (image_reid_inference.py:291-301).
    scikit-learn is a dependency of the reference (requirements.txt:11, unpinned; 1.9.0 in this image) and is not
    vendored; this restates its published algorithm: cluster/_dbscan.py (neighbourhoods = radius_neighbors(radius=eps)
    on the dense matrix = np.where(row <= eps), the python float eps compared in the row's float32; core = size >=
    min_samples, self included) and cluster/_dbscan_inner.pyx (for i ascending: an unlabelled core point seeds a new
    label and a depth-first search labels every unlabelled neighbour, pushing the core ones).
    Returns (labels int64[N], core bool[N])."""
    d = np.asarray(dists)
    N = d.shape[0]
    radius = d.dtype.type(eps)
    neigh = [np.where(d[i] <= radius)[0] for i in range(N)]
    core = np.array([len(n) >= min_samples for n in neigh], dtype=bool)
    labels = np.full(N, -1, dtype=np.int64)
    label_num = 0
    for i in range(N):
        if labels[i] != -1 or not core[i]:
            continue
        stack = []
        v = i
        while True:
            if labels[v] == -1:
                labels[v] = label_num
                if core[v]:
                    for w in neigh[v]:
                        if labels[w] == -1:
                            stack.append(int(w))
            if not stack:
                break
            v = stack.pop()
        label_num += 1
    return labels, core


def smooth_tracklets(embeddings: np.ndarray, seqs: np.ndarray, indices_valid: np.ndarray) -> np.ndarray:
    """reid/inference_utils.py:19-27: per unique sequence id, the valid rows become row * 0.1 + mean(valid rows) * 0.9
    (float32); a sequence id without valid rows is left alone (the reference's bare except)."""
    e = np.array(embeddings, dtype=np.float32, copy=True)
    seqs = np.asarray(seqs)
    valid = np.asarray(indices_valid, dtype=bool)
    for j in np.unique(seqs):
        idx = np.logical_and(seqs == j, valid)
        if not idx.any():
            continue
        cur = e[idx]
        avg = (cur.sum(0, dtype=np.float32) / np.float32(cur.shape[0])).astype(np.float32)
        e[idx] = cur * np.float32(0.1) + avg[None, :] * np.float32(0.9)
    return e


def diminish_camera_bias(embeddings: np.ndarray, cams: np.ndarray, la: float = 0.05) -> np.ndarray:
    """reid/inference_utils.py:5-16: per camera i in 0..max(cams): the rows of the camera lose their mean, are multiplied
    by P^T with P = inverse(X^T X + n * la * I) built from the rows BEFORE centring (the reference takes its copy
    first), and are renormalised to unit length.  float32 like the reference; the inverse in float64 then rounded."""
    e = np.array(embeddings, dtype=np.float32, copy=True)
    cams = np.asarray(cams).astype(np.int64)
    D = e.shape[1]
    for i in range(int(cams.max()) + 1):
        m = cams == i
        cur = e[m].copy()
        n = cur.shape[0]
        bias = (cur.sum(0, dtype=np.float32) / np.float32(n)).astype(np.float32) if n else np.zeros(D, np.float32)
        A = (cur.T @ cur + np.float32(n * la) * np.eye(D, dtype=np.float32)).astype(np.float32)
        P = np.linalg.inv(A.astype(np.float64)).astype(np.float32)
        y = ((cur - bias[None, :]) @ P.T).astype(np.float32)
        nrm = np.sqrt((y * y).sum(1, dtype=np.float32, keepdims=True))
        e[m] = y / nrm
    return e


# ----------------------------------------------------------------------------------------- scipy's assignment, ties included
def lsap_scipy_order(cost):
    """(row_ind, col_ind) exactly as scipy.optimize.linear_sum_assignment returns them, INCLUDING its choice between
    assignments of equal cost -- a statement-by-statement restatement of scipy/optimize/rectangular_lsap/
    rectangular_lsap.cpp (Crouse's shortest augmenting path; scipy is an un-vendored dependency of the reference:
    boxmot/trackers/strongsort/sort/linear_assignment.py:4,60).  The reference's unmatched lists inherit their order from
    these ties (every gated entry costs max_distance + 1e-5), so the device solver's exact mode follows this function;
    tests/test_oracle.py pins it against scipy itself on tie-heavy matrices."""
    cost = np.asarray(cost, dtype=np.float64)
    nr, nc = cost.shape
    tr = nc < nr
    if tr:
        cost = np.ascontiguousarray(cost.T)
        nr, nc = nc, nr
    u, v = np.zeros(nr), np.zeros(nc)
    col4row, row4col = -np.ones(nr, np.int64), -np.ones(nc, np.int64)
    for cur in range(nr):
        sp = np.full(nc, np.inf)
        path = -np.ones(nc, np.int64)
        SR, SC = np.zeros(nr, bool), np.zeros(nc, bool)
        remaining = [nc - it - 1 for it in range(nc)]
        num_remaining, sink, i, min_val = nc, -1, cur, 0.0
        while sink == -1:
            index, lowest = -1, np.inf
            SR[i] = True
            for it in range(num_remaining):
                j = remaining[it]
                r = min_val + cost[i, j] - u[i] - v[j]
                if r < sp[j]:
                    path[j] = i
                    sp[j] = r
                if sp[j] < lowest or (sp[j] == lowest and row4col[j] == -1):
                    lowest = sp[j]
                    index = it
            min_val = lowest
            if min_val == np.inf:
                raise ValueError("cost matrix is infeasible")
            j = remaining[index]
            if row4col[j] == -1:
                sink = j
            else:
                i = row4col[j]
            SC[j] = True
            num_remaining -= 1
            remaining[index] = remaining[num_remaining]
        u[cur] += min_val
        for i2 in range(nr):
            if SR[i2] and i2 != cur:
                u[i2] += min_val - sp[col4row[i2]]
        for j2 in range(nc):
            if SC[j2]:
                v[j2] -= min_val - sp[j2]
        j = sink
        while True:
            i2 = path[j]
            row4col[j] = i2
            col4row[i2], j = j, col4row[i2]
            if i2 == cur:
                break
    if tr:
        order = np.argsort(col4row)
        return col4row[order], np.arange(nr)[order]
    return np.arange(nr), col4row


def min_cost_matching_lists(cost, max_distance, track_indices, detection_indices, solver=None):
    """linear_assignment.py:52-77 on a ready cost matrix: (matches, unmatched_tracks, unmatched_detections) in the
    reference's order.  solver = scipy's linear_sum_assignment by default, lsap_scipy_order for the restatement."""
    if solver is None:
        from scipy.optimize import linear_sum_assignment as solver
    if len(detection_indices) == 0 or len(track_indices) == 0:
        return [], list(track_indices), list(detection_indices)
    c = np.array(cost, dtype=np.float64, copy=True)
    c[c > max_distance] = max_distance + 1e-5
    rows, cols = solver(c)
    matches, um_t, um_d = [], [], []
    for col, d in enumerate(detection_indices):
        if col not in cols:
            um_d.append(d)
    for row, t in enumerate(track_indices):
        if row not in rows:
            um_t.append(t)
    for row, col in zip(rows, cols):
        if c[row, col] > max_distance:
            um_t.append(track_indices[row])
            um_d.append(detection_indices[col])
        else:
            matches.append((track_indices[row], detection_indices[col]))
    return matches, um_t, um_d
