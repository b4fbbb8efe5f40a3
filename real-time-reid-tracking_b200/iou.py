"""Drop-in for yolov8_tracking/boxmot/utils/iou.py (iou_batch :4-23, giou_batch :26-59, diou_batch :62-102,
ciou_batch :105-164, get_asso_func :167-175): [N,4] x [K,4] xyxy boxes -> [N,K] numpy matrix in the inputs' dtype,
one kernel launch per call (csrc/boxpairs.cu).  The `*_device` variants return the CUDA tensor without the copy back."""
from __future__ import annotations

import numpy as np

from . import ops


def _run(b1, b2, mode):
    b1, b2 = np.asarray(b1), np.asarray(b2)
    dtype = np.result_type(b1.dtype, b2.dtype)
    dtype = np.float32 if dtype == np.float32 else np.float64
    if mode == "giou" and len(b1) and len(b2):
        # the reference asserts a non-degenerate enclosing box for every pair (iou.py:54)
        w = np.maximum(b1[:, None, 2], b2[None, :, 2]) - np.minimum(b1[:, None, 0], b2[None, :, 0])
        h = np.maximum(b1[:, None, 3], b2[None, :, 3]) - np.minimum(b1[:, None, 1], b2[None, :, 1])
        assert (w > 0).all() and (h > 0).all()
    return ops.box_pairs(b1.reshape(-1, 4), b2.reshape(-1, 4), mode, dtype)


def iou_batch(bboxes1, bboxes2):
    return _run(bboxes1, bboxes2, "iou").cpu().numpy()


def giou_batch(bboxes1, bboxes2):
    return _run(bboxes1, bboxes2, "giou").cpu().numpy()


def diou_batch(bboxes1, bboxes2):
    return _run(bboxes1, bboxes2, "diou").cpu().numpy()


def ciou_batch(bboxes1, bboxes2):
    return _run(bboxes1, bboxes2, "ciou").cpu().numpy()


def get_asso_func(asso_mode):
    ASSO_FUNCS = {"iou": iou_batch, "giou": giou_batch, "ciou": ciou_batch, "diou": diou_batch}
    return ASSO_FUNCS[asso_mode]
