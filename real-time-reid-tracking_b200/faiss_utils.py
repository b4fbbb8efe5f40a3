"""Drop-in for reid/faiss_utils.py: brute-force k-NN search helpers (:30-139), k_reciprocal_neigh (:142-146),
compute_jaccard_distance (:149-244) -- plus the classic re_ranking(q_g, q_q, g_g, k1, k2, lambda) entry the
north star asks for (not in the reference).

faiss itself is not used: its exact squared-L2 k-NN (bfKnn / GpuIndexFlatL2 / IndexFlatL2 / IndexShards)
is the tcgen05 GEMM + row-select of libreidb200, every `search_option` maps onto it.
"""
from __future__ import annotations

import time

import numpy as np
import torch

from . import _lib as L
from . import ops

METRIC_INNER_PRODUCT = 0  # faiss.METRIC_INNER_PRODUCT
METRIC_L2 = 1             # faiss.METRIC_L2


# ---- raw pointer helpers (reid/faiss_utils.py:8-27): same contiguity / dtype assertions, plain int address
def swig_ptr_from_FloatTensor(x):
    assert x.is_contiguous()
    assert x.dtype == torch.float32
    return x.data_ptr()


def swig_ptr_from_LongTensor(x):
    assert x.is_contiguous()
    assert x.dtype == torch.int64, 'dtype=%s' % x.dtype
    return x.data_ptr()


def swig_ptr_from_IntTensor(x):
    assert x.is_contiguous()
    assert x.dtype == torch.int32, 'dtype=%s' % x.dtype
    return x.data_ptr()


def _metric_name(metric):
    if metric == METRIC_L2:
        return "sql2"
    if metric == METRIC_INNER_PRODUCT:
        return "dot"
    raise ValueError("unsupported metric %r" % (metric,))


def search_raw_array_pytorch(res, xb, xq, k, D=None, I=None, metric=METRIC_L2):
    """bfKnn on torch tensors (faiss_utils.py:56-118): D f32[nq,k] (squared L2, ascending), I i32[nq,k].
    `res` (faiss GpuResources) is accepted and ignored."""
    assert xb.device == xq.device
    nq, d = xq.size()
    if not (xq.is_contiguous() or xq.t().is_contiguous()):
        raise TypeError('matrix should be row or column-major')
    nb, d2 = xb.size()
    assert d2 == d
    if not (xb.is_contiguous() or xb.t().is_contiguous()):
        raise TypeError('matrix should be row or column-major')
    if D is not None:
        assert D.shape == (nq, k)
        assert D.device == xb.device
    if I is not None:
        assert I.shape == (nq, k)
        assert I.device == xb.device
    Dd, Id = ops.knn(xq.contiguous(), xb.contiguous(), k, _metric_name(metric), idx64=False)
    if D is None:
        D = Dd.to(xb.device) if not xb.is_cuda else Dd
    else:
        D.copy_(Dd)
    if I is None:
        I = Id.to(xb.device) if not xb.is_cuda else Id
    else:
        I.copy_(Id)
    return D, I


def search_index_pytorch(index, x, k, D=None, I=None):
    """index.search with torch I/O (faiss_utils.py:30-53): D f32[n,k], I i64[n,k]."""
    assert x.is_contiguous()
    n, d = x.size()
    assert d == index.d
    if D is not None:
        assert D.size() == (n, k)
    if I is not None:
        assert I.size() == (n, k)
    Dd, Id = index.search_device(x, k)
    torch.cuda.synchronize()
    if D is None:
        D = Dd if x.is_cuda else Dd.cpu()
    else:
        D.copy_(Dd)
    if I is None:
        I = Id if x.is_cuda else Id.cpu()
    else:
        I.copy_(Id)
    return D, I


class FlatL2Index:
    """IndexFlatL2 / GpuIndexFlatL2 stand-in: vectors live on one GPU, exact squared-L2 search."""

    def __init__(self, d, device=None):
        self.d = int(d)
        self.device = device or L.require_device()
        self._xb = torch.empty((0, self.d), dtype=torch.float32, device=self.device)

    @property
    def ntotal(self):
        return self._xb.shape[0]

    def add(self, x):
        x = L.as_f32_cuda(x, self.device)
        assert x.shape[1] == self.d
        self._xb = torch.cat([self._xb, x], 0)

    def reset(self):
        self._xb = torch.empty((0, self.d), dtype=torch.float32, device=self.device)

    def search_device(self, x, k, index_base=0):
        return ops.knn(L.as_f32_cuda(x, self.device), self._xb, k, "sql2", index_base=index_base, idx64=True)

    def search(self, x, k):
        D, I = self.search_device(x, k)
        return D.cpu().numpy(), I.cpu().numpy()


class ShardedFlatL2Index:
    """faiss.IndexShards over one GpuIndexFlatL2 per GPU (faiss_utils.py:121-135), driven from ONE process:
    every add() splits its rows contiguously over the `ngpus` devices and may be called repeatedly (ids keep counting up
    in call order, as IndexShards(successive_ids) numbers them); every shard runs the fused GEMM + top-k on its own
    stream, the per-shard (distance, global id) lists are merged by the row-select kernel on device 0.
    (The one-process-per-GPU variant with an NCCL all-gather is reidb200.shard.ShardedSearcher.)"""

    def __init__(self, ngpus, d):
        avail = torch.cuda.device_count()
        if avail == 0:
            L.require_device()
        self.ngpus = max(1, min(int(ngpus) if ngpus else avail, avail))
        self.d = int(d)
        self.shards = [FlatL2Index(d, torch.device("cuda", i)) for i in range(self.ngpus)]
        self._ids = [None] * self.ngpus  # global id of every row a shard holds (i64, on the shard's device)
        self._n = 0

    @property
    def ntotal(self):
        return self._n

    def reset(self):
        for s in self.shards:
            s.reset()
        self._ids = [None] * self.ngpus
        self._n = 0

    def add(self, x):
        x = np.asarray(x, np.float32) if not isinstance(x, torch.Tensor) else x
        assert x.shape[1] == self.d
        n = x.shape[0]
        per = (n + self.ngpus - 1) // self.ngpus
        for i, s in enumerate(self.shards):
            lo, hi = min(n, i * per), min(n, (i + 1) * per)
            if hi > lo:
                s.add(x[lo:hi])
                ids = torch.arange(self._n + lo, self._n + hi, dtype=torch.int64, device=s.device)
                self._ids[i] = ids if self._ids[i] is None else torch.cat([self._ids[i], ids])
        self._n += n

    def search_device(self, x, k):
        parts_d, parts_i = [], []
        dev0 = self.shards[0].device
        for s, ids in zip(self.shards, self._ids):
            if s.ntotal == 0:
                continue
            with torch.cuda.device(s.device):
                D, I = s.search_device(x, k)
                valid = I >= 0  # fewer rows than k on this shard: faiss pads with -1
                I = torch.where(valid, ids[I.clamp(min=0)], I)
            parts_d.append(D.to(dev0, non_blocking=True))
            parts_i.append(I.to(dev0, non_blocking=True))
        for s in self.shards:
            torch.cuda.synchronize(s.device)
        if not parts_d:
            raise ValueError("search on an empty index")
        if len(parts_d) == 1:
            return parts_d[0], parts_i[0]
        with torch.cuda.device(dev0):
            return ops.topk_rows(torch.cat(parts_d, 1), k, descending=False, cand_i=torch.cat(parts_i, 1), idx64=True)

    def search(self, x, k):
        D, I = self.search_device(x, k)
        return D.cpu().numpy(), I.cpu().numpy()


def index_init_gpu(ngpus, feat_dim):
    index = ShardedFlatL2Index(ngpus, feat_dim)
    index.reset()
    return index


def index_init_cpu(feat_dim):
    """The reference builds a host IndexFlatL2; here the exact same search runs on the GPU (no CPU path)."""
    return FlatL2Index(feat_dim)


_krecip_cache = {}


def k_reciprocal_neigh(initial_rank, i, k1):
    """R(i,k1) (faiss_utils.py:142-146), from the device k-reciprocal kernel (lists for all rows are
    computed once per (table, k1) and cached)."""
    key = (id(initial_rank), int(k1), getattr(initial_rank, "shape", None))
    hit = _krecip_cache.get(key)
    if hit is None or hit[0] is not initial_rank:
        idx, cnt = ops.k_reciprocal_lists(torch.as_tensor(np.ascontiguousarray(np.asarray(initial_rank))), int(k1))
        _krecip_cache.clear()
        hit = (initial_rank, idx.cpu().numpy(), cnt.cpu().numpy())
        _krecip_cache[key] = hit
    _, idx, cnt = hit
    return idx[i, :cnt[i]].astype(np.asarray(initial_rank).dtype)


def compute_jaccard_distance_device(target_features, k1=20, k2=6, use_float16=False):
    """Device-resident result (torch CUDA tensor [N,N])."""
    return ops.jaccard(target_features, k1, k2, use_float16)


def compute_jaccard_distance(target_features, k1=20, k2=6, print_flag=True, search_option=0, use_float16=False):
    """np.ndarray[N,N] (float32, or float16 with use_float16) Jaccard distance of one merged set.
    search_option 0 / 1 / 3 (faiss GPU raw array / GPU index / CPU index): one GPU, the whole step in libreidb200.
    search_option 2 (faiss_utils.py:172-176: `index_init_gpu(ngpus, d)` = IndexShards over every visible GPU): the k-NN
    stage is sharded --
      * under torch.distributed (one process per GPU): shard.compute_jaccard_distance_sharded, rows of the merged set split
        over the ranks, one all-gather of the ranking tables, the N x N output row-sharded and gathered;
      * in a single process: the reference's own sequence, `index_init_gpu(ngpus, d)`, `index.add(features)`,
        `index.search(features, k1)` on the IndexShards stand-in (one shard per visible GPU, merged on device 0), then the
        sparse stages from that ranking (no N x N squared-L2 matrix is built)."""
    end = time.time()
    if print_flag:
        print('Computing jaccard distance...')
    if search_option == 2:
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            from . import shard

            J = shard.compute_jaccard_distance_sharded(target_features, k1, k2, use_float16)
        else:
            feats = L.as_f32_cuda(target_features)
            N = feats.shape[0]
            if not 1 <= k1 < N:
                raise ValueError("jaccard: need 1 <= k1 < N (k1=%d, N=%d)" % (k1, N))
            index = index_init_gpu(torch.cuda.device_count(), feats.shape[1])
            index.add(feats)
            D, initial_rank = index.search_device(feats, k1)
            J = ops.rerank_rows_from_rank(feats, initial_rank.to(feats.device), D.to(feats.device), k1, k2, (0, N), (0, N), 0.0, None,
                                          use_float16)
    else:
        J = ops.jaccard(target_features, k1, k2, use_float16)
    out = J.cpu().numpy()
    if print_flag:
        print("Jaccard distance computing time cost: {}".format(time.time() - end))
    return out


def re_ranking_feats(qf, gf, k1=20, k2=6, lambda_value=0.3, device_out=False):
    """Re-ranked Q x G distance from features: merged = cat(gallery, query) as image_reid_inference.py:270,
    Jaccard on the merged set, blended (1-lambda) * J + lambda * euclidean(q, g)."""
    qf = L.as_f32_cuda(qf)
    gf = L.as_f32_cuda(gf, qf.device)
    G, Q = gf.shape[0], qf.shape[0]
    merged = torch.cat([gf, qf], 0)
    out = ops.jaccard(merged, k1, k2, False, lambda_value, rows=(G, Q), cols=(0, G))
    return out if device_out else out.cpu().numpy()


def re_ranking(q_g_dist, q_q_dist, g_g_dist, k1=20, k2=6, lambda_value=0.3):
    """Classic k-reciprocal re-ranking signature (Zhong et al.), from Euclidean distance matrices of
    L2-normalised features: final = (1-lambda) * jaccard + lambda * q_g_dist, np.ndarray[Q,G].
    The Jaccard part is the reference's compute_jaccard_distance on the merged (gallery, query) set."""
    qg = L.as_f32_cuda(q_g_dist)
    dev = qg.device
    qq = L.as_f32_cuda(q_q_dist, dev)
    gg = L.as_f32_cuda(g_g_dist, dev)
    Q, G = qg.shape
    top = torch.cat([gg, qg.t()], 1)
    bot = torch.cat([qg, qq], 1)
    sql2 = torch.cat([top, bot], 0)
    sql2 = sql2 * sql2  # 2 - 2 x.y for unit vectors
    out = ops.jaccard_from_sql2(sql2, k1, k2, False, lambda_value, rows=(G, Q), cols=(0, G))
    return out.cpu().numpy()
