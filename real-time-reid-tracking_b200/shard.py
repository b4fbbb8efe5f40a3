"""Gallery-sharded exact k-NN across the GPUs of one box, one process per GPU.

Mirrors the reference's `index_init_gpu` (reid/faiss_utils.py:121-135): a faiss `IndexShards` over one
`GpuIndexFlatL2` per GPU -- contiguous row ranges per shard, queries replicated, per-shard top-k merged.  Here every
rank owns one shard as a RESIDENT packed operand (scaled fp16 hi|lo planes + squared norms, packed once at add time; the fp32
rows are not kept), runs the tcgen05 GEMM + chunk-guided row-select of libreidb200 on it (`reidb200_knn_packed`), and
the per-shard (distance, global id)[Q,k] tables are exchanged with ONE `all_gather` (NCCL over NVLink/NVSwitch,
Q*k*12 bytes per rank) and merged P*k -> k by the same row-select kernel with explicit candidate ids
(`reidb200_topk_rows`).  Ties resolve to the lower global id on every rank, so all ranks return identical tables.

`local_search` / `merge` are injection points for the host-logic tests (gloo, CPU tensors); the defaults are the CUDA
kernels and raise without an sm_100 device -- there is no CPU fallback.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

from . import _lib as L
from . import ops


def shard_bounds(n: int, world: int):
    """Contiguous split used by add(): shard r holds rows [lo_r, hi_r), ceil(n/world) rows each (the last ones may be
    short or empty)."""
    per = (n + world - 1) // world if world > 0 else n
    return [(min(n, r * per), min(n, (r + 1) * per)) for r in range(world)]


class ShardedSearcher:
    def __init__(self, d, group=None, device=None, local_search=None, merge=None):
        self.d = int(d)
        self.group = group
        self.distributed = dist.is_available() and dist.is_initialized()
        self.rank = dist.get_rank(group) if self.distributed else 0
        self.world = dist.get_world_size(group) if self.distributed else 1
        self._custom = local_search is not None
        self._local_search = local_search
        self._merge = merge
        if self._custom:
            self.device = torch.device("cpu") if device is None else torch.device(device)
        else:
            self.device = device or L.require_device()
        self.reset()

    # ------------------------------------------------------------------ shard contents
    def reset(self):
        self.base = 0          # global id of local row 0
        self.n_local = 0
        self.capacity = 0
        self._packed = None    # u8 [capacity * packed_row_bytes]
        self._sqnorm = None    # f32 [capacity]
        self._iscale = None    # f32 [capacity] inverse power-of-two row scales of the packed planes
        self._rows = None      # injected-search mode keeps the rows as given

    @property
    def ntotal(self):
        """Rows over all shards (collective when distributed)."""
        if not self.distributed:
            return self.n_local
        t = torch.tensor([self.n_local], dtype=torch.int64, device=self.device)
        dist.all_reduce(t, group=self.group)
        return int(t.item())

    def reserve(self, rows):
        """Preallocate the packed shard so a large gallery can be appended piece by piece (no 2x peak)."""
        if self._custom:
            return
        rows = int(rows)
        rb = ops.packed_row_bytes(self.d)
        packed = torch.empty(max(rows * rb, 256), dtype=torch.uint8, device=self.device)
        sq = torch.empty(max(rows, 1), dtype=torch.float32, device=self.device)
        isc = torch.empty(max(rows, 1), dtype=torch.float32, device=self.device)
        if self.n_local:
            packed[: self.n_local * rb].copy_(self._packed[: self.n_local * rb])
            sq[: self.n_local].copy_(self._sqnorm[: self.n_local])
            isc[: self.n_local].copy_(self._iscale[: self.n_local])
        self._packed, self._sqnorm, self._iscale, self.capacity = packed, sq, isc, rows

    def append(self, x):
        """Append local rows f32[n,d] (host or device) to this rank's shard; global ids continue from base."""
        n = int(x.shape[0])
        if n == 0:
            return
        if x.shape[1] != self.d:
            raise ValueError("append: feature dimension %d != %d" % (x.shape[1], self.d))
        if self._custom:
            x = torch.as_tensor(np.asarray(x, np.float32)) if not isinstance(x, torch.Tensor) else x.float()
            self._rows = x if self._rows is None else torch.cat([self._rows, x], 0)
            self.n_local += n
            return
        if self.n_local + n > self.capacity:
            self.reserve(max(self.n_local + n, 2 * self.capacity))
        rb = ops.packed_row_bytes(self.d)
        lo = self.n_local
        ops.pack_rows(L.as_f32_cuda(x, self.device), out=self._packed[lo * rb:(lo + n) * rb], sqnorm=self._sqnorm[lo:lo + n],
                      iscale=self._iscale[lo:lo + n])
        self.n_local += n

    def add_shard(self, x_local, base=None):
        """This rank's rows; `base` = global id of its first row (default: exclusive prefix sum of the shard sizes over
        ranks, i.e. shards are numbered in rank order like IndexShards(successive_ids=True))."""
        if self.n_local:
            raise NotImplementedError("add_shard once after reset (contiguous shards); use append() to extend")
        self.append(x_local)
        if base is None:
            base = 0
            if self.distributed:
                sizes = torch.zeros(self.world, dtype=torch.int64, device=self.device)
                sizes[self.rank] = self.n_local
                dist.all_reduce(sizes, group=self.group)
                base = int(sizes[: self.rank].sum().item())
        self.base = int(base)

    def add(self, x):
        """Whole gallery given on every rank (faiss `index.add(x)` on an IndexShards): keep only this rank's slice."""
        lo, hi = shard_bounds(int(x.shape[0]), self.world)[self.rank]
        self.add_shard(x[lo:hi], base=lo)

    # ------------------------------------------------------------------ search
    def _search_local(self, xq, k, metric):
        if self._custom:
            return self._local_search(xq, self._rows, k, metric, self.base)
        return ops.knn_packed(xq, self._packed, self._sqnorm, self._iscale, self.n_local, self.d, k, metric, index_base=self.base, idx64=True)

    def search_device(self, xq, k, metric="sql2"):
        """(D f32[Q,k], I i64[Q,k]) identical on every rank; queries must be the same on every rank."""
        if not self._custom:
            xq = L.as_f32_cuda(xq, self.device)
        nq = int(xq.shape[0])
        pad = float("-inf") if metric == "dot" else float("inf")
        if self.n_local:
            Dl, Il = self._search_local(xq, k, metric)
        else:
            Dl = torch.full((nq, k), pad, dtype=torch.float32, device=self.device)
            Il = torch.full((nq, k), -1, dtype=torch.int64, device=self.device)
        if self.world == 1:
            return Dl, Il
        Dg = torch.empty((self.world * nq, k), dtype=torch.float32, device=self.device)
        Ig = torch.empty((self.world * nq, k), dtype=torch.int64, device=self.device)
        dist.all_gather_into_tensor(Dg, Dl.contiguous(), group=self.group)
        dist.all_gather_into_tensor(Ig, Il.contiguous(), group=self.group)
        cand_d = Dg.view(self.world, nq, k).permute(1, 0, 2).reshape(nq, self.world * k).contiguous()
        cand_i = Ig.view(self.world, nq, k).permute(1, 0, 2).reshape(nq, self.world * k).contiguous()
        if self._custom:
            return self._merge(cand_d, cand_i, k, metric)
        return ops.topk_rows(cand_d, k, descending=(metric == "dot"), cand_i=cand_i, idx64=True)

    def search(self, xq, k, metric="sql2"):
        """faiss-style host results (D f32[Q,k], I i64[Q,k]) as numpy."""
        D, I = self.search_device(xq, k, metric)
        return D.cpu().numpy(), I.cpu().numpy()


# --------------------------------------------------------------------------------------------- query-sharded CMC / mAP
def _eval_queries_cuda(qf, ql, qc, gf, gl, gc):
    """Cumulative first-hit counts f64[G] and AP sum f64[] of a query slice on this rank's GPU: tcgen05 score GEMM + the
    counting kernel of csrc/evalmap.cu (a query with no valid match adds nothing, like evaluate.py:41-43).  The kernel
    returns float(count) / float(q); count <= 2^24 so round(cmc * q) restores the integer exactly -- no extra pass, no
    host sync."""
    score = ops.distmat(qf, gf, "dot")
    cmc, _, ap_q, _ = ops.eval_cmc_map(score, ql, qc, gl, gc, descending=True, per_query=True)
    return torch.round(cmc.to(torch.float64) * float(qf.shape[0])), ap_q.sum()


def evaluate_all_sharded(qf, ql, qc, gf, gl, gc, group=None, device=None, local_eval=None):
    """`evaluate_all` (reid/evaluate.py:33-52) with the QUERIES split over the ranks (queries are
    independent).  Every rank holds the gallery, evaluates queries [lo_r, hi_r) of `shard_bounds(Q, world)` and
    contributes (#queries whose first hit is at rank <= r, for every r) and its sum of APs; ONE all-reduce(sum) of
    f64[G+1] (counts are exact in f64) gives every rank the same `(CMC f32[G], mAP f64)`, divided by ALL Q queries as
    the reference does.  Counts are integers, so CMC is bit-identical to the single-GPU call; the AP sum differs from
    it by the summation order only (<= 1e-15).

    `local_eval(qf, ql, qc, gf, gl, gc) -> (ap_q f64[q], first i32[q])` is the injection point of the gloo host-logic
    test; the default is the CUDA path and raises without an sm_100 device."""
    distributed = dist.is_available() and dist.is_initialized()
    rank = dist.get_rank(group) if distributed else 0
    world = dist.get_world_size(group) if distributed else 1
    custom = local_eval is not None
    if custom:
        dev = torch.device("cpu") if device is None else torch.device(device)
        as_f = lambda x: torch.as_tensor(np.asarray(x, np.float32) if not isinstance(x, torch.Tensor) else x).to(dev)
    else:
        dev = device or L.require_device()
        as_f = lambda x: L.as_f32_cuda(x, dev)
    lab = lambda x: (x if isinstance(x, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(np.asarray(x)))).reshape(-1)
    qf, gf = as_f(qf), as_f(gf)
    ql, qc, gl, gc = lab(ql), lab(qc), lab(gl), lab(gc)
    Q, G = int(qf.shape[0]), int(gf.shape[0])
    if Q == 0 or G == 0:
        raise ValueError("evaluate_all: empty query or gallery set")
    if ql.numel() != Q or qc.numel() != Q or gl.numel() != G or gc.numel() != G:
        raise ValueError("label / camera vectors do not match the feature matrices")
    lo, hi = shard_bounds(Q, world)[rank]
    part = torch.zeros(G + 1, dtype=torch.float64, device=dev)
    if hi > lo and custom:
        ap_q, first = local_eval(qf[lo:hi], ql[lo:hi], qc[lo:hi], gf, gl, gc)
        first = first.to(dev).long()
        hist = torch.bincount(first[first >= 0], minlength=G)[:G]
        part[:G] = torch.cumsum(hist, 0).to(torch.float64)
        part[G] = ap_q.to(dev).sum()
    elif hi > lo:
        part[:G], part[G] = _eval_queries_cuda(qf[lo:hi], ql[lo:hi], qc[lo:hi], gf, gl, gc)
    if world > 1:
        dist.all_reduce(part, op=dist.ReduceOp.SUM, group=group)
    # tensor / tensor: an IEEE division like the finalize kernel's (torch turns `/ python_scalar` into `* (1/Q)`)
    cmc = part[:G].to(torch.float32) / torch.full((1,), float(Q), dtype=torch.float32, device=dev)
    ap = part[G] / Q
    return cmc, ap


# --------------------------------------------------------------------------------------------- row-sharded re-ranking
def _gather_rows(local, n_total, world, group):
    """all-gather of a row-sharded table (rank r holds rows shard_bounds(n_total, world)[r], all but the last shards have
    ceil(n_total / world) rows): ONE all_gather_into_tensor of equal-sized (padded) pieces; returns [n_total, ...]."""
    if world == 1:
        return local
    per = (n_total + world - 1) // world
    piece = local
    if piece.shape[0] != per:
        pad = torch.zeros((per - piece.shape[0],) + tuple(piece.shape[1:]), dtype=piece.dtype, device=piece.device)
        piece = torch.cat([piece, pad], 0)
    out = torch.empty((world * per,) + tuple(piece.shape[1:]), dtype=piece.dtype, device=piece.device)
    dist.all_gather_into_tensor(out, piece.contiguous(), group=group)
    return out[:n_total]


def initial_rank_sharded(feats, k1, group=None, local_knn=None):
    """Stage D2 of the re-ranking split over the ranks (the reference's search_option 2 builds the same table
    through faiss IndexShards, reid/faiss_utils.py:172-176): rank r searches its ROWS [lo_r, hi_r) of the merged set against
    all N rows (tcgen05 GEMM + row select) and the [rows, k1] pieces are exchanged with one all-gather each for the indices
    and the squared-L2 values.  Returns (rank_i i32[N,k1], rank_d f32[N,k1]), identical on every rank."""
    distributed = dist.is_available() and dist.is_initialized()
    rank = dist.get_rank(group) if distributed else 0
    world = dist.get_world_size(group) if distributed else 1
    N = int(feats.shape[0])
    lo, hi = shard_bounds(N, world)[rank]
    if local_knn is not None:
        D, I = local_knn(feats[lo:hi], feats, k1)
    else:
        D, I = ops.knn(feats[lo:hi], feats, k1, "sql2", idx64=False)
    rank_i = _gather_rows(I.to(torch.int32), N, world, group)
    rank_d = _gather_rows(D.to(torch.float32), N, world, group)
    return rank_i, rank_d


def rerank_eval_sharded(qf, ql, qc, gf, gl, gc, k1=20, k2=6, lambda_value=0.3, group=None, device=None, local_knn=None,
                        local_rows=None):
    """ONE k-reciprocal re-ranked evaluation across the ranks of `group` (one process per GPU), SURVEY 8e row 3:

      1. k-NN stage, row-sharded: every rank searches N / P rows of merged = cat(gallery, query) against all N rows;
         all-gather of the (index, squared L2)[N, k1] tables (2 x N k1 4 bytes in total: 3 MB at Market size).
      2. sparse stages (reciprocal lists, expansion, query expansion, inverted index) from the global table: replicated --
         they are ~0.25 ms of latency-bound kernels on 0.5 GB of tables, cheaper to redo than to exchange.
      3. Jaccard + lambda blend + CMC / mAP counting for this rank's Q / P QUERY rows only (its block of the Q x G squared-L2
         matrix comes from one more tcgen05 GEMM of Q / P x G), then ONE all-reduce(sum) of f64[G + 1] (first-hit counts +
         AP sum), as evaluate_all_sharded does.

    Returns (CMC f32[G], mAP f64[]) identical on every rank and equal to ops.rerank_eval on one GPU (same kernels on the same
    pair values: CMC bit-identical, the AP sum up to summation order).  `local_knn` / `local_rows` are the injection points
    of the gloo host-logic test."""
    distributed = dist.is_available() and dist.is_initialized()
    rank = dist.get_rank(group) if distributed else 0
    world = dist.get_world_size(group) if distributed else 1
    custom = local_knn is not None
    if custom:
        dev = torch.device("cpu") if device is None else torch.device(device)
        as_f = lambda x: torch.as_tensor(np.asarray(x, np.float32) if not isinstance(x, torch.Tensor) else x).to(dev)
    else:
        dev = device or L.require_device()
        as_f = lambda x: L.as_f32_cuda(x, dev)
    lab = lambda x: (x if isinstance(x, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(np.asarray(x)))).reshape(-1).to(dev).to(torch.int32)
    qf, gf = as_f(qf), as_f(gf)
    ql, qc, gl, gc = lab(ql), lab(qc), lab(gl), lab(gc)
    Q, G = int(qf.shape[0]), int(gf.shape[0])
    if Q == 0 or G == 0:
        raise ValueError("rerank_eval_sharded: empty query or gallery set")
    merged = torch.cat([gf, qf], 0)
    rank_i, rank_d = initial_rank_sharded(merged, k1, group, local_knn)
    q0, q1 = shard_bounds(Q, world)[rank]
    part = torch.zeros(G + 1, dtype=torch.float64, device=dev)
    if q1 > q0:
        if custom:
            hist, ap_q = local_rows(merged, rank_i, rank_d, k1, k2, G + q0, q1 - q0, G, lambda_value, ql[q0:q1], qc[q0:q1], gl, gc)
        else:
            blend = None
            if lambda_value > 0:
                ld = (G + 31) // 32 * 32  # 128-byte rows: the fused kernel reads the block with 128-bit loads
                blend = torch.empty((q1 - q0, ld), dtype=torch.float32, device=dev)
                ops.distmat(merged[G + q0:G + q1], gf, "sql2", out=blend[:, :G])
            _, hist, ap_q, _ = ops.rerank_rows_from_rank(merged, rank_i, rank_d, k1, k2, (G + q0, q1 - q0), (0, G), lambda_value, blend,
                                                         want_out=False, labels=(ql[q0:q1], qc[q0:q1], gl, gc))
        part[:G] = torch.cumsum(hist.to(torch.float64), 0)
        part[G] = ap_q.sum()
    if world > 1:
        dist.all_reduce(part, op=dist.ReduceOp.SUM, group=group)
    cmc = part[:G].to(torch.float32) / torch.full((1,), float(Q), dtype=torch.float32, device=dev)
    return cmc, part[G] / Q


def compute_jaccard_distance_sharded(feats, k1=20, k2=6, use_float16=False, group=None, gather=True):
    """compute_jaccard_distance (reid/faiss_utils.py:149-244) across the ranks of `group`: the k-NN stage row-sharded with
    one all-gather of the ranking tables, the sparse stages replicated, and the dense N x N OUTPUT row-sharded -- rank r
    writes rows shard_bounds(N, world)[r].  gather=True all-gathers the row blocks so that every rank returns the full
    matrix (the reference's contract); gather=False returns (row block, (lo, hi)) and never holds N x N on one GPU."""
    distributed = dist.is_available() and dist.is_initialized()
    rank = dist.get_rank(group) if distributed else 0
    world = dist.get_world_size(group) if distributed else 1
    feats = L.as_f32_cuda(feats)
    N = int(feats.shape[0])
    rank_i, rank_d = initial_rank_sharded(feats, k1, group)
    lo, hi = shard_bounds(N, world)[rank]
    block = ops.rerank_rows_from_rank(feats, rank_i, rank_d, k1, k2, (lo, hi - lo), (0, N), 0.0, None, use_float16)
    if not gather:
        return block, (lo, hi)
    return _gather_rows(block, N, world, group)
