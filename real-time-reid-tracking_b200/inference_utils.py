"""Drop-in for reid/inference_utils.py: the post-embedding tricks of image_reid_inference.py on device.
Same names, arguments and in-place behaviour as the reference; inputs may be host or CUDA tensors (host tensors are
copied to the device, processed there and written back)."""
from __future__ import annotations

import numpy as np
import torch

from . import _lib as L
from . import ops


def _to_dev(embeddings):
    dev = embeddings.device if embeddings.is_cuda else L.require_device()
    e = embeddings if (embeddings.is_cuda and embeddings.dtype == torch.float32 and embeddings.is_contiguous()) \
        else embeddings.to(device=dev, dtype=torch.float32).contiguous()
    return e, dev


def diminish_camera_bias(embeddings, cams, la=0.05):
    """reid/inference_utils.py:5-16.  Per camera i = 0..max(cams): the camera's rows lose their mean, are multiplied by
    P^T with P = inverse(X^T X + n * la * I) (X = the rows before centring) and are renormalised.  The Gram matrix and
    the projection run through the fp32-accurate tcgen05 GEMM (ops.distmat, kind "dot"), centring and normalisation
    through reidb200 kernels; the D x D inverse is torch.linalg.inv on the device (the call the reference makes)."""
    e, dev = _to_dev(embeddings)
    cams_d = torch.as_tensor(cams).to(dev)
    num_cams = int(cams_d.max().item())
    D = e.shape[1]
    for i in range(num_cams + 1):
        idx = torch.nonzero(cams_d == i).reshape(-1)
        n = int(idx.numel())
        cur = e.index_select(0, idx)
        if n == 0:  # the reference's inverse of the zero matrix raises; mirror it
            raise RuntimeError("diminish_camera_bias: camera %d has no rows (singular ridge system)" % i)
        cur_t = cur.t().contiguous()
        A = ops.distmat(cur_t, cur_t, "dot")
        A.diagonal().add_(n * la)
        P = torch.linalg.inv(A)
        y = ops.distmat(ops.center_columns(cur), P, "dot")  # (cur - mean) @ P.t()
        e.index_copy_(0, idx, ops.tta_merge_normalize(y))
    if e.data_ptr() != embeddings.data_ptr():
        embeddings.copy_(e)
    return embeddings


def smooth_tracklets(embeddings, seqs, indices_valid):
    """reid/inference_utils.py:19-27: per sequence id, the valid rows become row * 0.1 + mean(valid rows) * 0.9."""
    e, dev = _to_dev(embeddings)
    seqs_d = torch.as_tensor(seqs).to(dev).reshape(-1)
    valid = torch.as_tensor(np.asarray(indices_valid.cpu() if isinstance(indices_valid, torch.Tensor) else indices_valid)
                            .astype(bool)).to(dev).reshape(-1)
    rows = torch.nonzero(valid).reshape(-1)
    if rows.numel():
        keys, perm = torch.sort(seqs_d[rows], stable=True)
        _, counts = torch.unique_consecutive(keys, return_counts=True)
        seg = torch.zeros(counts.numel() + 1, dtype=torch.int32, device=dev)
        seg[1:] = torch.cumsum(counts, 0).to(torch.int32)
        ops.smooth_groups_(e, rows[perm].to(torch.int32), seg, 0.1, 0.9)
    if e.data_ptr() != embeddings.data_ptr():
        embeddings.copy_(e)
    return embeddings


def cluster_smooth_evaluate(gallery_embeddings, query_embeddings, gallery_labels, query_labels, gallery_cams, query_cams,
                            gallery_seqs, query_seqs, eps=0.5, min_samples=7, attribute_dist=None, camera_bias=True,
                            k1=20, k2=6):
    """The experimental inference path of reid/image_reid_inference.py:246-315 as one device-resident pipeline:
    flip-TTA merge of the (plain, flipped) embedding pairs -> cat(gallery, query) -> diminish_camera_bias ->
    compute_jaccard_distance -> clamp + attribute distance -> DBSCAN -> smooth_tracklets -> evaluate_all.
    `*_embeddings` = a pair (plain, flipped) or a single already merged tensor.  Returns (CMC FloatTensor[G], mAP float,
    pseudo_labels i64 ndarray); the N x N matrix never leaves the device."""
    from . import evaluate

    dev = L.require_device()

    def merge(e):
        if isinstance(e, (tuple, list)):
            return ops.tta_merge_normalize(torch.as_tensor(e[0]).to(dev), torch.as_tensor(e[1]).to(dev))
        return torch.as_tensor(e).to(device=dev, dtype=torch.float32)

    g, q = merge(gallery_embeddings), merge(query_embeddings)
    G = g.shape[0]
    merged = torch.cat((g, q), dim=0)
    cat = lambda a, b: torch.cat((torch.as_tensor(a).reshape(-1), torch.as_tensor(b).reshape(-1))).to(dev)
    merged_cams, merged_seqs = cat(gallery_cams, query_cams), cat(gallery_seqs, query_seqs)
    if camera_bias:
        merged = diminish_camera_bias(merged, merged_cams)
    dists = ops.jaccard(merged, k1, k2)
    ops.clamp_add_(dists, attribute_dist)
    pseudo, ncl, _ = ops.dbscan_precomputed(dists, eps, min_samples)
    del dists
    num_labels = int(ncl.item())
    merged_seqs = merged_seqs.to(torch.int64) * num_labels + pseudo.to(torch.int64)  # image_reid_inference.py:305
    merged = smooth_tracklets(merged, merged_seqs, pseudo != -1)
    CMC, ap = evaluate.evaluate_all(merged[G:], query_labels, query_cams, merged[:G], gallery_labels, gallery_cams)
    return CMC, ap, pseudo.cpu().numpy().astype(np.int64)
