"""Drop-in for modification_deepsort/feature_extractor.py (Extractor :11-53) and the preprocessing half of
boxmot's ReIDDetectMultiBackend (yolov8_tracking/boxmot/appearance/reid_multibackend.py:210-220).

Only the crop -> resize -> normalise step is a libreidb200 kernel; the embedding backbone stays a PyTorch
module supplied by the caller (the reference's own import is broken, SURVEY F10, so `net` is injectable and a
random-init stand-in with a 512-d output is used when none is given)."""
from __future__ import annotations

import numpy as np
import torch

from . import preprocess as P


class _StandInBackbone(torch.nn.Module):
    """Random-init 512-d embedding (stride-8 conv stem + global pool): backbone weights are out of scope."""

    def __init__(self, dim=512):
        super().__init__()
        self.stem = torch.nn.Sequential(torch.nn.Conv2d(3, 32, 7, 4, 3), torch.nn.ReLU(), torch.nn.Conv2d(32, 64, 3, 2, 1),
                                        torch.nn.ReLU(), torch.nn.AdaptiveAvgPool2d(1))
        self.fc = torch.nn.Linear(64, dim)

    def forward(self, x):
        return self.fc(self.stem(x).flatten(1))


class Extractor(object):
    def __init__(self, model_path=None, use_cuda=True, net=None, standin=False):
        """`net`: the embedding backbone (a torch module).  The reference builds SEDense18_IBN from a module that does not
        exist in its tree, so the backbone has to come from the caller; `model_path` (a checkpoint with a
        "net_dict" or a plain state dict) is loaded into it STRICTLY.  Without `net` the call raises -- silently returning
        embeddings of random weights is not a drop-in -- unless standin=True asks for the random-init 512-d stand-in
        (tests / benchmarks only: backbone weights are outside the kernel scope)."""
        self.device = "cuda" if use_cuda else "cpu"
        if not use_cuda or not torch.cuda.is_available():
            raise RuntimeError("reidb200 Extractor needs a CUDA device (no CPU fallback)")
        if net is None:
            if not standin:
                raise ValueError("Extractor: pass the embedding backbone as `net` (the reference's SEDense18_IBN import is "
                                 "broken, SURVEY F10); standin=True builds a random-init stand-in for tests")
            if model_path is not None:
                raise ValueError("Extractor: a checkpoint cannot be loaded into the random-init stand-in backbone")
            net = _StandInBackbone()
        self.net = net
        if model_path is not None:
            state = torch.load(model_path, map_location="cpu")
            self.net.load_state_dict(state.get("net_dict", state) if isinstance(state, dict) else state, strict=True)
        self.net.to(self.device).eval()
        self.size = (128, 256)                       # (w, h), feature_extractor.py:24
        self.norm = P.DEEPSORT_NORM                  # Normalize([.5]*3, [.5]*3), :27-29

    def _preprocess(self, im_crops):
        """list of HWC uint8 crops -> float32 [M,3,256,128] (on the CUDA device):
        u8/255 -> cv2-style bilinear resize to (128,256) -> (x-0.5)/0.5 (feature_extractor.py:31-46)."""
        return P.preprocess_crops(im_crops, self.size, self.norm, "cv_linear")

    def preprocess_frame(self, frame, xyxy):
        """B200-first entry: the frame is uploaded (or already resident) once, crops are taken on the device."""
        return P.preprocess_frame(frame, xyxy, self.size, self.norm, "cv_linear")

    def __call__(self, im_crops):
        im_batch = self._preprocess(im_crops)
        with torch.no_grad():
            features = self.net(im_batch)
        return features.cpu().numpy()


class ReIDPreprocess(object):
    """ReIDDetectMultiBackend._preprocess: ToPILImage -> Resize((256,128)) antialiased bilinear on uint8 ->
    ToTensor -> Normalize(ImageNet) -> stack -> device (reid_multibackend.py:58-66,210-220)."""

    def __init__(self, image_size=(256, 128)):
        self.image_size = image_size  # (h, w) as torchvision Resize takes it

    def _preprocess(self, im_batch):
        return P.preprocess_crops(im_batch, (self.image_size[1], self.image_size[0]), P.IMAGENET_NORM, "pil_antialias")

    def preprocess_frame(self, frame, xyxy):
        return P.preprocess_frame(frame, xyxy, (self.image_size[1], self.image_size[0]), P.IMAGENET_NORM, "pil_antialias")
