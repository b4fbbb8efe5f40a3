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
        """labels i32[N], core 0/1 i32[N], n_clusters i32[1] as CUDA tensors, no host sync.  X must be symmetric (fit()
        checks a sample of pairs; this entry does not, to stay sync-free)."""
        if isinstance(X, torch.Tensor) and X.is_cuda:
            d = X if X.dtype == torch.float32 else X.float()
        else:
            d = torch.as_tensor(np.asarray(X, dtype=np.float32)).to(L.require_device())
        if d.dim() != 2 or d.shape[0] != d.shape[1]:
            raise ValueError("Precomputed distance matrix must be square")
        if d.stride(1) != 1:
            d = d.contiguous()
        return ops.dbscan_precomputed(d, self.eps, self.min_samples)

    @staticmethod
    def _check_symmetric(d, samples=8192):
        """The union pass visits every core pair once through the lower triangle, i.e. it assumes d[i, j] == d[j, i] (true
        for the Jaccard matrix and for dist + attribute_dist of image_reid_inference.py:284-301).  sklearn's definition on
        an asymmetric matrix is row-neighbourhood based and would differ silently, so a sampled check refuses such input."""
        N = d.shape[0]
        if N < 2:
            return
        g = torch.Generator(device="cpu").manual_seed(N)
        i = torch.randint(0, N, (samples,), generator=g).to(d.device)
        j = torch.randint(0, N, (samples,), generator=g).to(d.device)
        if float((d[i, j] - d[j, i]).abs().max()) > 1e-5:  # (rounding-level asymmetry of a GEMM-produced matrix is fine)
            raise ValueError("reidb200.cluster.DBSCAN needs a SYMMETRIC precomputed distance matrix (d[i, j] == d[j, i])")

    def fit(self, X, y=None, sample_weight=None):
        if sample_weight is not None:
            raise ValueError("sample_weight is not supported")
        if isinstance(X, torch.Tensor) and X.is_cuda:
            self._check_symmetric(X)
        else:
            Xh = np.asarray(X)
            if Xh.ndim == 2 and Xh.shape[0] == Xh.shape[1] and not np.allclose(Xh, Xh.T, rtol=0, atol=1e-5):
                raise ValueError("reidb200.cluster.DBSCAN needs a SYMMETRIC precomputed distance matrix (d[i, j] == d[j, i])")
        labels, _, core = self.fit_device(X)
        self.labels_ = labels.cpu().numpy().astype(np.int64)
        self.core_sample_indices_ = np.nonzero(core.cpu().numpy())[0]
        return self

    def fit_predict(self, X, y=None, sample_weight=None):
        return self.fit(X, sample_weight=sample_weight).labels_
