"""The slice of sklearn.cluster.DBSCAN that reid/image_reid_inference.py:291-301 uses, on the device-resident matrix:
DBSCAN(eps, min_samples, metric="precomputed").fit_predict(dists).  `dists` may be the CUDA tensor
compute_jaccard_distance_device returns (no 1.49 GB device->host copy) or a host array (copied to the device)."""
from __future__ import annotations

import numpy as np
import torch

from . import _lib as L
from . import ops


class DBSCAN:
    def __init__(self, eps=0.5, min_samples=5, metric="precomputed", n_jobs=None, **unused):
        if metric != "precomputed":
            raise ValueError("reidb200.cluster.DBSCAN implements metric='precomputed' only (the reference's call)")
        if not eps > 0.0:
            raise ValueError("eps must be positive")
        self.eps, self.min_samples, self.metric, self.n_jobs = float(eps), int(min_samples), metric, n_jobs

    def fit_device(self, X):
        """labels i32[N], core 0/1 i32[N], n_clusters i32[1] as CUDA tensors, no host sync."""
        if isinstance(X, torch.Tensor) and X.is_cuda:
            d = X if X.dtype == torch.float32 else X.float()
        else:
            d = torch.as_tensor(np.asarray(X, dtype=np.float32)).to(L.require_device())
        if d.dim() != 2 or d.shape[0] != d.shape[1]:
            raise ValueError("Precomputed distance matrix must be square")
        if d.stride(1) != 1:
            d = d.contiguous()
        return ops.dbscan_precomputed(d, self.eps, self.min_samples)

    def fit(self, X, y=None, sample_weight=None):
        if sample_weight is not None:
            raise ValueError("sample_weight is not supported")
        labels, _, core = self.fit_device(X)
        self.labels_ = labels.cpu().numpy().astype(np.int64)
        self.core_sample_indices_ = np.nonzero(core.cpu().numpy())[0]
        return self

    def fit_predict(self, X, y=None, sample_weight=None):
        return self.fit(X, sample_weight=sample_weight).labels_
