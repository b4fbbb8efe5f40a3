"""Seeded synthetic inputs shaped like the reference's workloads.

Host-side numpy only.  Market1501-shaped query/gallery embeddings with labels and
cameras (configs C1/C2), MOT16/MOT20-shaped frames, boxes and tracker state (C3/C4).
"""
from __future__ import annotations

import numpy as np


def _unit(x):
    return (x / np.linalg.norm(x, axis=1, keepdims=True)).astype(np.float32)


def hard_noise(D):
    """Noise level at which a same-id pair sits about 3 sigma above the different-id cosines of D-dimensional
    features (2.5 at D = 512): plain rank-1 ~ 0.9-0.95, mAP ~ 0.5-0.85, first hits spread over many ranks --
    Market-like difficulty instead of the trivially separable default (noise = 0.6: every AP is 0 or 1)."""
    return float(np.sqrt(max(np.sqrt(float(D)) / 3.1 - 1.0, 0.25)))


def market_like(Q=3368, G=15913, D=512, n_ids=751, n_cams=6, seed=0, noise=0.6,
                junk_frac=0.18, orphan_frac=0.01):
    """Returns dict(qf,ql,qc,gf,gl,gc): fp32 L2-normalised features, int64 labels/cams.
    noise = 0.6 is SURVEY 8d's recipe (easy: rank-1 == mAP); noise = "hard" uses hard_noise(D).

    Gallery: ids 0..n_ids-1 plus label -1 distractors (junk_frac of G).  Every query has at
    least one cross-camera gallery match, except ~orphan_frac of them whose id is absent
    from the gallery (exercises the -1 sentinel of reid/evaluate.py:80-82).
    """
    if noise == "hard":
        noise = hard_noise(D)
    rng = np.random.default_rng(seed)
    n_ids = max(2, min(n_ids, max(2, G // 4)))
    cent = rng.standard_normal((n_ids + 1, D)).astype(np.float32)
    n_junk = int(G * junk_frac)
    gl = np.concatenate([np.arange(G - n_junk) % n_ids, -np.ones(n_junk, np.int64)]).astype(np.int64)
    rng.shuffle(gl)
    gc = rng.integers(0, n_cams, G).astype(np.int64)
    base = np.where(gl[:, None] >= 0, cent[np.maximum(gl, 0)], rng.standard_normal((G, D)).astype(np.float32))
    gf = _unit(base + noise * rng.standard_normal((G, D)).astype(np.float32))
    ql = rng.integers(0, n_ids, Q).astype(np.int64)
    qc = rng.integers(0, n_cams, Q).astype(np.int64)
    orphan = rng.random(Q) < orphan_frac
    ql[orphan] = n_ids  # id that never occurs in the gallery
    qf = _unit(cent[ql] + noise * rng.standard_normal((Q, D)).astype(np.float32))
    # guarantee a cross-camera positive for non-orphans: flip the camera of one gallery item
    for i in np.where(~orphan)[0][: min(Q, 4096)]:
        same = np.where(gl == ql[i])[0]
        if same.size and not np.any(gc[same] != qc[i]):
            gc[same[0]] = (qc[i] + 1) % n_cams
    return dict(qf=qf, ql=ql, qc=qc, gf=gf, gl=gl, gc=gc)


def mot_frame(M=30, T=100, budget=100, D=512, H=1080, W=1920, seed=0, fresh_frac=0.1):
    """One tracking frame: image, detections and per-track state.

    Returns dict with
      img u8[H,W,3]; det_xywh f32[M,4] (centre); det_tlwh f32[M,4]; det_feat f32[M,D];
      samples: list of T arrays f32[count_t,D] (unit rows, EMA-like successive states);
      trk_mean f64[T,8] (xyah + velocities); trk_SI f64[T,4,4] (zeros for "fresh" tracks,
      SURVEY F6); trk_tsu i32[T] time_since_update.
    """
    rng = np.random.default_rng(seed)
    img = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    h = rng.uniform(60, 400, M)
    w = h * rng.uniform(0.3, 0.5, M)
    cx = rng.uniform(-20, W + 20, M)
    cy = rng.uniform(-20, H + 20, M)
    det_xywh = np.stack([cx, cy, w, h], 1).astype(np.float32)
    det_tlwh = det_xywh.copy()
    det_tlwh[:, 0] -= det_xywh[:, 2] / 2
    det_tlwh[:, 1] -= det_xywh[:, 3] / 2
    det_feat = rng.standard_normal((M, D)).astype(np.float32)
    samples, means, SIs = [], [], []
    for t in range(T):
        cnt = int(rng.integers(1, budget + 1))
        # about half of the tracks look like one of the detections
        anchor = det_feat[t % M] if t < M or rng.random() < 0.5 else rng.standard_normal(D).astype(np.float32)
        cur = anchor / np.linalg.norm(anchor)
        rows = []
        for _ in range(cnt):
            new = cur + 0.15 * rng.standard_normal(D).astype(np.float32)
            cur = 0.9 * cur + 0.1 * new / np.linalg.norm(new)
            cur = cur / np.linalg.norm(cur)
            rows.append(cur.astype(np.float32))
        samples.append(np.stack(rows, 0))
        m = t % M
        hh = float(det_xywh[m, 3]) * rng.uniform(0.9, 1.1)
        mean = np.zeros(8)
        mean[0] = det_xywh[m, 0] + rng.normal(0, 6)
        mean[1] = det_xywh[m, 1] + rng.normal(0, 6)
        mean[2] = float(det_xywh[m, 2] / det_xywh[m, 3]) * rng.uniform(0.9, 1.1)
        mean[3] = hh
        means.append(mean)
        if rng.random() < fresh_frac:
            SIs.append(np.zeros((4, 4)))
        else:
            std = np.array([hh / 20, hh / 20, 1e-2, hh / 20]) * rng.uniform(0.5, 3.0)
            A = rng.standard_normal((4, 4)) * 0.1
            S = np.diag(std ** 2) + (A @ A.T) * np.outer(std, std)
            SIs.append(np.linalg.inv(S))
    tsu = rng.choice(np.array([1, 1, 1, 2, 5], np.int32), T)
    return dict(img=img, det_xywh=det_xywh, det_tlwh=det_tlwh, det_feat=det_feat, samples=samples,
                trk_mean=np.stack(means, 0), trk_SI=np.stack(SIs, 0), trk_tsu=tsu.astype(np.int32))
