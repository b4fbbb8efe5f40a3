"""Drop-in for reid/evaluate.py (evaluate_all :33-52, evaluate :55-75, compute_mAP :78-105, get_feats :5-20,
get_evaluate_result :23-27).

The Q x G score matrix is one tensor-core GEMM (reid/evaluate.py:58 does Q separate GEMVs + Q host
syncs + Q full sorts); CMC / mAP come from the counting kernel in csrc/evalmap.cu -- no sort, one
device->host copy of G+1 numbers at the end.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib as L
from . import ops


def _labels(x):
    if isinstance(x, torch.Tensor):
        return x.reshape(-1)
    return torch.from_numpy(np.ascontiguousarray(np.asarray(x))).reshape(-1)


def _check_ap(ap, what):
    """The counting kernels hold at most EV_GOOD_CAP = 1024 good matches per query (csrc/evalmap.cuh); a query with more
    poisons the AP sum with NaN (and sets the device status word).  Every host-returning entry point turns that into an error."""
    if ap != ap:
        raise L.Reidb200Error("%s: a query has more good gallery matches than this build holds (1024)" % what)
    return ap


def evaluate_all_device(qf, ql, qc, gf, gl, gc):
    """Device-resident variant: returns (CMC f32[G], mAP f64[1]) CUDA tensors, no host sync."""
    qf = L.as_f32_cuda(qf)
    gf = L.as_f32_cuda(gf, qf.device)
    if qf.shape[0] == 0 or gf.shape[0] == 0:
        raise ValueError("evaluate_all: empty query or gallery set")
    score = ops.distmat(qf, gf, "dot")
    return ops.eval_cmc_map(score, _labels(ql), _labels(qc), _labels(gl), _labels(gc), descending=True)


def evaluate_all(qf, ql, qc, gf, gl, gc):
    """(CMC FloatTensor[G] averaged over ALL Q queries, mAP float); prints Rank@1/5/10 + mAP like the reference."""
    cmc, ap = evaluate_all_device(qf, ql, qc, gf, gl, gc)
    CMC = cmc.cpu()
    ap = _check_ap(float(ap.cpu()[0]), "evaluate_all")
    r = lambda k: float(CMC[k]) if CMC.numel() > k else float("nan")
    print('Rank@1:%f Rank@5:%f Rank@10:%f mAP:%f' % (r(0), r(4), r(9), ap))
    return CMC, ap


def evaluate_all_sharded(qf, ql, qc, gf, gl, gc, group=None, verbose=True):
    """`evaluate_all` with the queries split over the ranks of `group` (one process per GPU, gallery replicated) and
    ONE NCCL all-reduce of the CMC counts + AP sum (shard.evaluate_all_sharded); same return value on every rank."""
    from . import shard

    cmc, ap = shard.evaluate_all_sharded(qf, ql, qc, gf, gl, gc, group=group)
    CMC, ap = cmc.cpu(), _check_ap(float(ap.cpu()), "evaluate_all_sharded")
    if verbose and shard.dist.is_initialized() and shard.dist.get_rank(group) != 0:
        verbose = False
    if verbose:
        r = lambda k: float(CMC[k]) if CMC.numel() > k else float("nan")
        print('Rank@1:%f Rank@5:%f Rank@10:%f mAP:%f' % (r(0), r(4), r(9), ap))
    return CMC, ap


def evaluate_dist(dist, ql, qc, gl, gc):
    """Same reduction on a precomputed Q x G DISTANCE matrix (ascending = better), e.g. a re-ranked one."""
    cmc, ap = ops.eval_cmc_map(dist, _labels(ql), _labels(qc), _labels(gl), _labels(gc), descending=False)
    return cmc.cpu(), _check_ap(float(ap.cpu()[0]), "evaluate_dist")


def evaluate(qf, ql, qc, gf, gl, gc):
    """One query: (ap float, cmc IntTensor[G]); cmc[0] == -1 when the query has no valid gallery match."""
    qf = L.as_f32_cuda(qf).reshape(1, -1)
    gf = L.as_f32_cuda(gf, qf.device)
    G = gf.shape[0]
    score = ops.distmat(qf, gf, "dot")
    one = lambda v: torch.as_tensor(v).reshape(1)
    _, _, ap_q, first = ops.eval_cmc_map(score, one(ql), one(qc), _labels(gl), _labels(gc), descending=True, per_query=True)
    first = int(first.cpu()[0])
    cmc = torch.IntTensor(G).zero_()
    if first < 0:
        cmc[0] = -1
        return 0, cmc
    cmc[first:] = 1
    return float(ap_q.cpu()[0]), cmc


def compute_mAP(index, good_index, junk_index):
    """(ap, cmc IntTensor[len(index)]) for a given ranking `index` (best first) -- runs the same CUDA
    counting kernel on a synthetic score row (score = -position)."""
    index = np.asarray(index).reshape(-1)
    good_index = np.asarray(good_index).reshape(-1)
    junk_index = np.asarray(junk_index).reshape(-1)
    n = len(index)
    cmc = torch.IntTensor(n).zero_()
    if good_index.size == 0:
        cmc[0] = -1
        return 0, cmc
    G = int(max(index.max(initial=-1), good_index.max(initial=-1), junk_index.max(initial=-1))) + 1
    score = np.full((1, G), -float(n + 1), np.float32)
    score[0, index] = -np.arange(n, dtype=np.float32)
    gl = np.zeros(G, np.int32)
    gc = np.zeros(G, np.int32)
    gl[junk_index.astype(np.int64)] = -1
    gl[good_index.astype(np.int64)] = 1
    gc[good_index.astype(np.int64)] = 1
    not_ranked = np.ones(G, bool)
    not_ranked[index] = False
    gl[not_ranked & (gl == 0)] = -1  # items absent from `index` never count
    _, _, ap_q, first = ops.eval_cmc_map(torch.from_numpy(score), np.array([1], np.int32), np.array([0], np.int32),
                                         gl, gc, descending=True, per_query=True)
    first = int(first.cpu()[0])
    cmc[first:] = 1
    return float(ap_q.cpu()[0]), cmc


def get_feats(model, dataloader):
    """Backbone forward over a loader (PyTorch; the backbone is outside the kernel scope)."""
    model.eval()
    labels, features, cams = [], [], []
    for sample in dataloader:
        image, label, cam = sample[:3]
        feat = model(image.cuda(), cam)
        features.append(feat[1])
        labels.append(label)
        cams.append(cam)
    labels = torch.cat(labels, dim=0)
    features = torch.cat(features, dim=0)
    cams = torch.cat(cams, dim=0)
    model.train()
    return labels, features, cams


def get_evaluate_result(model, dataloader_query, dataloader_gallery):
    labels_query, features_query, cams_query = get_feats(model, dataloader_query)
    labels_gallery, features_gallery, cams_gallery = get_feats(model, dataloader_gallery)
    return evaluate_all(features_query, labels_query, cams_query, features_gallery, labels_gallery, cams_gallery)


def _merge_into(merged, gf, qf):
    """merged[:G] = gf, merged[G:] = qf without a torch.cat temporary: host inputs are copied straight into place."""
    G = gf.shape[0]
    for dst, src in ((merged[:G], gf), (merged[G:], qf)):
        if not isinstance(src, torch.Tensor):
            src = torch.from_numpy(np.ascontiguousarray(np.asarray(src, dtype=np.float32)))
        dst.copy_(src, non_blocking=True)
    return merged


_plan = {}  # the most recent shape's captured evaluation (RerankEvalPlan), reused by back-to-back calls


def release_plans():
    """Drop the cached CUDA graph and its buffers (workspace of about 4 * N^2 bytes)."""
    _plan.clear()


def _copy_labels(lab, Q, G, ql, qc, gl, gc):
    off, views = 0, []
    for v, n in ((ql, Q), (qc, Q), (gl, G), (gc, G)):  # the order of RerankEvalPlan.lab
        t = _labels(v)
        if t.numel() != n:
            raise ValueError("label / camera vectors do not match the query / gallery sets")
        if t.dtype != torch.int32:
            t = t.to(torch.int32)
        lab[off:off + n].copy_(t, non_blocking=True)
        views.append(lab[off:off + n])
        off += n
    return views


def evaluate_all_reranked_device(qf, ql, qc, gf, gl, gc, k1=20, k2=6, lambda_value=0.3, merged=None):
    """Re-ranked evaluation, device resident: merged = cat(gallery, query) (image_reid_inference.py:270), Jaccard
    re-ranking blended with the Euclidean distance, CMC/mAP on the re-ranked Q x G distance.  Returns CUDA tensors
    (CMC f32[G], mAP f64[1]) without a host sync.  With `merged` (a preallocated f32[G+Q, D] device buffer) the kernels
    are launched one by one; without it the call replays the CUDA graph cached for its shape (ops.RerankEvalPlan)."""
    G, Q, D = int(gf.shape[0]), int(qf.shape[0]), int(qf.shape[1])
    dev = qf.device if isinstance(qf, torch.Tensor) and qf.is_cuda else L.require_device()
    if merged is None and 0 < G <= ops.RERANK_EVAL_MAX_G and Q > 0:
        key = (dev, G, Q, D, int(k1), int(k2), float(lambda_value))
        plan = _plan.get(key)
        if plan is None:
            _plan.clear()
            with torch.cuda.device(dev):
                plan = _plan[key] = ops.RerankEvalPlan(G, Q, D, k1, k2, lambda_value, dev)
        with torch.cuda.device(dev):
            _merge_into(plan.merged, gf, qf)
            _copy_labels(plan.lab, Q, G, ql, qc, gl, gc)
            cmc, ap = plan.run()
            return cmc.clone(), ap.clone()
    if merged is None:
        merged = torch.empty((G + Q, D), dtype=torch.float32, device=dev)
    _merge_into(merged, gf, qf)
    return ops.rerank_eval(merged, G, _labels(ql), _labels(qc), _labels(gl), _labels(gc), k1, k2, lambda_value)


class RerankedEvalStream:
    """Back-to-back re-ranked evaluations of many (query, gallery) sets with the copies hidden: the host->device copy
    of set i+1 runs on a copy stream while set i is computed, results come back through pinned buffers.

        ev = RerankedEvalStream(depth=2)
        t0 = ev.submit(qf, ql, qc, gf, gl, gc)        # returns at once
        t1 = ev.submit(...)                            # its H2D overlaps the kernels of t0
        CMC, ap = ev.result(t0)                        # same values as evaluate_all_reranked

    One evaluation is the unit the reference computes in image_reid_inference.py:270-315; this class only schedules
    several of them (CUDA streams + events, no threads)."""

    def __init__(self, depth=2, device=None):
        self.dev = device or L.require_device()
        self.depth = max(1, int(depth))
        self.copy_stream = torch.cuda.Stream(self.dev)
        self.compute_stream = torch.cuda.Stream(self.dev)
        self.slots = [dict(busy=None, free=None) for _ in range(self.depth)]
        self.n = 0

    def _buffers(self, slot, Q, G, D, k1, k2, lambda_value):
        key = (Q, G, D, int(k1), int(k2), float(lambda_value))
        if slot.get("key") != key:
            with torch.cuda.device(self.dev):
                # a captured CUDA graph per slot when the fused kernel covers the gallery, plain launches otherwise
                slot["plan"] = ops.RerankEvalPlan(G, Q, D, k1, k2, lambda_value, self.dev) if G <= ops.RERANK_EVAL_MAX_G else None
                if slot["plan"] is not None:
                    slot["merged"], slot["lab"] = slot["plan"].merged, slot["plan"].lab
                else:
                    slot["merged"] = torch.empty((G + Q, D), dtype=torch.float32, device=self.dev)
                    slot["lab"] = torch.empty(2 * (Q + G), dtype=torch.int32, device=self.dev)
                slot["cmc_h"] = torch.empty(G, dtype=torch.float32).pin_memory()
                slot["ap_h"] = torch.empty(1, dtype=torch.float64).pin_memory()
            slot["key"] = key
        return slot

    def submit(self, qf, ql, qc, gf, gl, gc, k1=20, k2=6, lambda_value=0.3):
        G, Q, D = int(gf.shape[0]), int(qf.shape[0]), int(qf.shape[1])
        ticket = self.n
        slot = self.slots[ticket % self.depth]
        if slot["busy"] is not None:  # the slot's previous evaluation must have been collected or at least finished
            slot["busy"].synchronize()
        self._buffers(slot, Q, G, D, k1, k2, lambda_value)
        lab = slot["lab"]
        with torch.cuda.stream(self.copy_stream):
            if slot["free"] is not None:
                self.copy_stream.wait_event(slot["free"])  # kernels of the slot's previous occupant are done
            _merge_into(slot["merged"], gf, qf)
            views = _copy_labels(lab, Q, G, ql, qc, gl, gc)
            ready = torch.cuda.Event()
            ready.record(self.copy_stream)
        with torch.cuda.stream(self.compute_stream):
            self.compute_stream.wait_event(ready)
            if slot["plan"] is not None:
                cmc, ap = slot["plan"].run()
            else:
                cmc, ap = ops.rerank_eval(slot["merged"], G, views[0], views[1], views[2], views[3], k1, k2, lambda_value)
            free = torch.cuda.Event()
            free.record(self.compute_stream)
            slot["cmc_h"].copy_(cmc, non_blocking=True)
            slot["ap_h"].copy_(ap, non_blocking=True)
            done = torch.cuda.Event()
            done.record(self.compute_stream)
        slot["free"], slot["busy"], slot["ticket"] = free, done, ticket
        self.n += 1
        return ticket

    def result(self, ticket):
        slot = self.slots[ticket % self.depth]
        if slot.get("ticket") != ticket:
            raise ValueError("RerankedEvalStream.result: ticket %d is no longer held (depth=%d)" % (ticket, self.depth))
        slot["busy"].synchronize()
        return slot["cmc_h"].clone(), _check_ap(float(slot["ap_h"][0]), "RerankedEvalStream.result")


def evaluate_all_reranked(qf, ql, qc, gf, gl, gc, k1=20, k2=6, lambda_value=0.3, verbose=True):
    """(CMC FloatTensor[G], mAP float) of the k-reciprocal re-ranked ranking; host arrays in, host results out."""
    cmc, ap = evaluate_all_reranked_device(qf, ql, qc, gf, gl, gc, k1, k2, lambda_value)
    CMC = cmc.cpu()
    ap = _check_ap(float(ap.cpu()[0]), "evaluate_all_reranked")
    if verbose:
        r = lambda k: float(CMC[k]) if CMC.numel() > k else float("nan")
        print('Rank@1:%f Rank@5:%f Rank@10:%f mAP:%f' % (r(0), r(4), r(9), ap))
    return CMC, ap
