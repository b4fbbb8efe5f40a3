"""Batched ROI crop + resize + normalise (north_star item 1), host-side geometry + one kernel launch.

  xywh_to_xyxy          StrongSORT._xywh_to_xyxy (boxmot/trackers/strongsort/strong_sort.py:115-121), vectorised in the
                        input's own float dtype exactly like the per-box Python arithmetic
  crops_from_frame      StrongSORT._get_features crop semantics (strong_sort.py:149-159): img[y1:y2, x1:x2]
  preprocess_frame      frame already on the device (or uploaded once) + boxes -> f32[M,3,H,W]
  preprocess_crops      list of host crops (the reference's _preprocess input) packed into one pinned buffer
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib as L
from . import ops

DEEPSORT_NORM = ((0.5, 0.5, 0.5), (0.5, 0.5, 0.5))                    # feature_extractor.py:27-29
IMAGENET_NORM = ((0.485, 0.456, 0.406), (0.229, 0.224, 0.225))        # reid_multibackend.py:58-66


def xywh_to_xyxy(boxes_xywh, width, height):
    """Centre-xywh [M,4] -> int32 [M,4] (x1,y1,x2,y2): x1 = max(int(x - w/2), 0), x2 = min(int(x + w/2), W-1), ...
    The arithmetic runs in the dtype of the input (float32 for YOLO boxes), int() truncates toward zero."""
    b = boxes_xywh.detach().cpu().numpy() if isinstance(boxes_xywh, torch.Tensor) else np.asarray(boxes_xywh)
    if b.dtype not in (np.float32, np.float64):
        b = b.astype(np.float64)
    b = b.reshape(-1, 4)
    half_w, half_h = b[:, 2] / 2, b[:, 3] / 2
    x1 = np.maximum(np.trunc(b[:, 0] - half_w).astype(np.int64), 0)
    x2 = np.minimum(np.trunc(b[:, 0] + half_w).astype(np.int64), width - 1)
    y1 = np.maximum(np.trunc(b[:, 1] - half_h).astype(np.int64), 0)
    y2 = np.minimum(np.trunc(b[:, 1] + half_h).astype(np.int64), height - 1)
    return np.stack([x1, y1, x2, y2], 1).astype(np.int32)


def _geometry_from_xyxy(xyxy, width, height=None):
    xyxy = np.asarray(xyxy, np.int64).reshape(-1, 4)
    if height is not None and len(xyxy):
        # numpy's img[y1:y2, x1:x2] clips silently at the frame border (negative starts aside); a box that leaves the
        # frame would make the kernel read neighbouring rows or past the buffer, so it is clipped the same way here
        xyxy = xyxy.copy()
        if (xyxy[:, :2] < 0).any():
            raise ValueError("crop box with a negative origin (StrongSORT._xywh_to_xyxy clamps at 0)")
        xyxy[:, 2] = np.minimum(xyxy[:, 2], width)
        xyxy[:, 3] = np.minimum(xyxy[:, 3], height)
    w = xyxy[:, 2] - xyxy[:, 0]
    h = xyxy[:, 3] - xyxy[:, 1]
    if len(xyxy) and (w.min() <= 0 or h.min() <= 0):
        # the reference hands an empty array to cv2.resize / PIL, which raises
        raise ValueError("empty crop: box %d has zero area after clamping" % int(np.argmin(np.minimum(w, h))))
    off = (xyxy[:, 1] * width + xyxy[:, 0]) * 3
    geom = np.stack([h, w, np.full_like(h, width * 3)], 1).astype(np.int32)
    return off.astype(np.int64), geom


def preprocess_frame(frame, xyxy, size=(128, 256), norm=DEEPSORT_NORM, mode="cv_linear", out=None):
    """frame u8[H,W,3] (torch CUDA tensor, or host array uploaded once), xyxy int [M,4] -> f32[M,3,size[1],size[0]] CUDA."""
    if isinstance(frame, torch.Tensor) and frame.is_cuda:
        dev = L.require_device(frame)
        f = frame
    else:
        dev = L.require_device()
        f = torch.as_tensor(np.ascontiguousarray(frame)).to(dev, non_blocking=True)
    if f.dtype != torch.uint8 or f.dim() != 3 or f.shape[2] != 3 or not f.is_contiguous():
        raise ValueError("frame must be a contiguous uint8 HxWx3 image")
    H, W = int(f.shape[0]), int(f.shape[1])
    off, geom = _geometry_from_xyxy(xyxy, W, H)
    if len(off) == 0:
        return torch.empty((0, 3, size[1], size[0]), dtype=torch.float32, device=dev)
    off_d, geom_d = L.staging(dev).upload([off, geom])  # one pinned H2D for both tables
    return ops.crop_resize_norm(f.reshape(-1), off_d, geom_d, size[1], size[0], norm[0], norm[1], mode, out=out)


def preprocess_crops(im_crops, size=(128, 256), norm=DEEPSORT_NORM, mode="cv_linear"):
    """List of host HWC uint8 crops (any sizes) -> f32[M,3,size[1],size[0]] CUDA: one packed upload, one launch."""
    dev = L.require_device()
    M = len(im_crops)
    if M == 0:
        return torch.empty((0, 3, size[1], size[0]), dtype=torch.float32, device=dev)
    geom = np.zeros((M, 3), np.int32)
    off = np.zeros(M, np.int64)
    total = 0
    for i, c in enumerate(im_crops):
        c = np.asarray(c)
        if c.ndim != 3 or c.shape[2] != 3 or c.shape[0] == 0 or c.shape[1] == 0:
            raise ValueError("crop %d has shape %s; HxWx3 with non-zero area expected" % (i, c.shape))
        geom[i] = (c.shape[0], c.shape[1], c.shape[1] * 3)
        off[i] = total
        total += c.shape[0] * c.shape[1] * 3
    packed = torch.empty(total, dtype=torch.uint8, pin_memory=True)
    pn = packed.numpy()
    for i, c in enumerate(im_crops):
        n = geom[i, 0] * geom[i, 1] * 3
        pn[off[i]:off[i] + n] = np.ascontiguousarray(c, dtype=np.uint8).reshape(-1)
    src = packed.to(dev, non_blocking=True)
    off_d = torch.from_numpy(off).to(dev, non_blocking=True)
    geom_d = torch.from_numpy(geom).to(dev, non_blocking=True)
    return ops.crop_resize_norm(src, off_d, geom_d, size[1], size[0], norm[0], norm[1], mode)


def resize_pil_u8_crops(im_crops, size=(128, 256)):
    """Bit-exact PIL.Image.resize((w,h), BILINEAR) of host crops, u8[M,h,w,3] CUDA (for parity tests)."""
    dev = L.require_device()
    M = len(im_crops)
    geom = np.array([(c.shape[0], c.shape[1], c.shape[1] * 3) for c in im_crops], np.int32).reshape(-1, 3)
    sizes = geom[:, 0].astype(np.int64) * geom[:, 1] * 3
    off = np.concatenate([[0], np.cumsum(sizes)[:-1]]).astype(np.int64) if M else np.zeros(0, np.int64)
    flat = np.concatenate([np.ascontiguousarray(c, dtype=np.uint8).reshape(-1) for c in im_crops]) if M else np.zeros(0, np.uint8)
    return ops.resize_pil_u8(torch.from_numpy(flat).to(dev), torch.from_numpy(off).to(dev), torch.from_numpy(geom).to(dev),
                             int(geom[:, 0].max()) if M else 1, size[1], size[0])
