"""Drop-in for the crop / feature half of yolov8_tracking/boxmot/trackers/strongsort/strong_sort.py: the box conversions
(:99-147) and StrongSORT._get_features(bbox_xywh, img) (:149-159).

The reference slices img[y1:y2, x1:x2] per box on the host and hands the list of crops to the ReID model, which resizes
and normalises them one by one (ReIDDetectMultiBackend._preprocess, reid_multibackend.py:210-220).  Here the frame is
uploaded once (or is already on the device) and ONE kernel launch produces the normalised batch f32[M,3,256,128] for all
boxes; the embedding backbone stays the caller's PyTorch module.  Track lifecycle (StrongSORT.update, :56-97) is host
Python outside the kernel scope; tracker.match_arrays / tracker.match are the association entry points."""
from __future__ import annotations

import numpy as np
import torch

from . import preprocess as P


class StrongSORT(object):
    """The members of the reference class that sit on the hot path.  `model` maps the preprocessed CUDA batch
    f32[M,3,h,w] to features (any torch module / callable); `preprocess_mode` selects the reference front-end:
    "pil_antialias" = boxmot's ReIDDetectMultiBackend (ImageNet normalisation), "cv_linear" = the DeepSort Extractor."""

    def __init__(self, model, max_dist=0.2, max_iou_distance=0.7, max_age=70, n_init=3, nn_budget=100, mc_lambda=0.995,
                 ema_alpha=0.9, preprocess_mode="pil_antialias", image_size=(256, 128)):
        from .matching import NearestNeighborDistanceMetric

        self.model = model
        self.max_dist = max_dist
        self.max_iou_distance, self.max_age, self.n_init = max_iou_distance, max_age, n_init
        self.mc_lambda, self.ema_alpha = mc_lambda, ema_alpha
        self.metric = NearestNeighborDistanceMetric("cosine", self.max_dist, nn_budget)   # strong_sort.py:32
        self.preprocess_mode = preprocess_mode
        self.image_size = image_size  # (h, w)
        self.height = self.width = None

    # ---- box conversions (strong_sort.py:99-147)
    @staticmethod
    def _xywh_to_tlwh(bbox_xywh):
        if isinstance(bbox_xywh, np.ndarray):
            bbox_tlwh = bbox_xywh.copy()
        elif isinstance(bbox_xywh, torch.Tensor):
            bbox_tlwh = bbox_xywh.clone()
        bbox_tlwh[:, 0] = bbox_xywh[:, 0] - bbox_xywh[:, 2] / 2.0
        bbox_tlwh[:, 1] = bbox_xywh[:, 1] - bbox_xywh[:, 3] / 2.0
        return bbox_tlwh

    def _xywh_to_xyxy(self, bbox_xywh):
        x1, y1, x2, y2 = P.xywh_to_xyxy(np.asarray(bbox_xywh).reshape(1, 4), self.width, self.height)[0]
        return int(x1), int(y1), int(x2), int(y2)

    def _tlwh_to_xyxy(self, bbox_tlwh):
        x, y, w, h = bbox_tlwh
        x1 = max(int(x), 0)
        x2 = min(int(x + w), self.width - 1)
        y1 = max(int(y), 0)
        y2 = min(int(y + h), self.height - 1)
        return x1, y1, x2, y2

    def _xyxy_to_tlwh(self, bbox_xyxy):
        x1, y1, x2, y2 = bbox_xyxy
        return x1, y1, int(x2 - x1), int(y2 - y1)

    # ---- crops -> features (strong_sort.py:149-159)
    def preprocess(self, bbox_xywh, img):
        """f32[M,3,h,w] CUDA batch of the boxes' crops img[y1:y2, x1:x2], resized and normalised in one launch."""
        frame = img if isinstance(img, torch.Tensor) else np.asarray(img)
        self.height, self.width = int(frame.shape[0]), int(frame.shape[1])
        boxes = bbox_xywh.detach().cpu().numpy() if isinstance(bbox_xywh, torch.Tensor) else np.asarray(bbox_xywh)
        xyxy = P.xywh_to_xyxy(boxes, self.width, self.height)
        norm = P.IMAGENET_NORM if self.preprocess_mode == "pil_antialias" else P.DEEPSORT_NORM
        return P.preprocess_frame(frame, xyxy, (self.image_size[1], self.image_size[0]), norm, self.preprocess_mode)

    @torch.no_grad()
    def _get_features(self, bbox_xywh, img):
        """features of every box (what self.model returns for the batch); np.array([]) when there is no box.  A box
        whose clamped crop is empty raises, as the reference's resize of an empty crop does."""
        if len(bbox_xywh) == 0:
            self.height, self.width = int(img.shape[0]), int(img.shape[1])
            return np.array([])
        return self.model(self.preprocess(bbox_xywh, img))
