"""torchrun --nproc-per-node N tests/nccl_shard_check.py : the NCCL path of the gallery-sharded search against (a) the same
search on ONE resident gallery (rank 0, bit-identical tables expected: same kernels, same tie rule) and (b) the CPU oracle
on a sub-problem.  Prints one line per check; exit code 0 = all equal."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

import reidb200  # noqa: F401
from oracle import reid_oracle as O
from reidb200 import ops, shard

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
ok = True
for nq, nb, d, k in ((64, 20_001, 128, 100), (300, 150_000, 256, 100), (17, 5, 32, 10)):
    rng = np.random.default_rng(nb)
    xb = rng.standard_normal((nb, d)).astype(np.float32)
    xb /= np.linalg.norm(xb, axis=1, keepdims=True)
    xq = rng.standard_normal((nq, d)).astype(np.float32)
    xq /= np.linalg.norm(xq, axis=1, keepdims=True)
    if nb > 100:
        xb[nb - 3] = xb[11]  # duplicate rows in the first and the last shard
        xq[0] = xb[11]
    s = shard.ShardedSearcher(d, device=dev)
    s.add(xb)
    D, I = s.search(xq, k)
    D1, I1 = ops.knn(xq, xb, k, "sql2")
    same = np.array_equal(I, I1.cpu().numpy()) and np.array_equal(D, D1.cpu().numpy())
    kk = min(k, nb)
    Do, Io = O.knn_l2(xb, xq[:16], kk)
    bad = I[:16, :kk] != Io
    tie_ok = True
    if bad.any():
        full = O.sql2_dist(xq[:16], xb).astype(np.float64)
        r, c = np.where(bad)
        tie_ok = bool(np.allclose(full[r, I[:16, :kk][r, c]], full[r, Io[r, c]], rtol=1e-4, atol=2e-5))
    dup_ok = nb <= 100 or (I[0, 0] == 11 and I[0, 1] == nb - 3)
    good = same and tie_ok and dup_ok and np.allclose(D[:16, :kk], Do, rtol=1e-4, atol=2e-5)
    ok &= bool(good)
    if rank == 0:
        print("shard_check world=%d nq=%d nb=%d d=%d k=%d: sharded==single %s, oracle %s, dup order %s" %
              (world, nq, nb, d, k, same, tie_ok, dup_ok), flush=True)
# query-sharded evaluate_all (one all-reduce of CMC counts + AP sum) against the single-GPU call
from reidb200 import evaluate, synth

for Q, G, D in ((3368, 15913, 512), (5, 40, 32), (1, 9, 16)):
    m = synth.market_like(Q=Q, G=G, D=D, seed=Q)
    a = [m[k] for k in ("qf", "ql", "qc", "gf", "gl", "gc")]
    cmc, ap = shard.evaluate_all_sharded(*a, device=dev)
    cmc1, ap1 = evaluate.evaluate_all_device(*[torch.from_numpy(x).to(dev) for x in a])
    same = bool(torch.equal(cmc, cmc1)) and abs(float(ap) - float(ap1[0])) < 1e-12
    ok &= same
    if rank == 0:
        print("shard_check world=%d evaluate_all Q=%d G=%d: query-sharded == single %s (rank1 %.6f mAP %.6f)" %
              (world, Q, G, same, float(cmc[0]), float(ap)), flush=True)
# row-sharded re-ranked evaluation (k-NN rows sharded + all-gather of the ranking tables, query rows sharded + all-reduce)
# against the single-GPU fused evaluation; and the row-sharded dense Jaccard matrix against the single-GPU one
from reidb200 import faiss_utils

for Q, G, D, hard in ((3368, 15913, 512, True), (300, 2100, 64, True), (3, 50, 32, False)):
    m = synth.market_like(Q=Q, G=G, D=D, seed=Q + 1, noise="hard" if hard else 0.6)
    a = [m[k] for k in ("qf", "ql", "qc", "gf", "gl", "gc")]
    k1, k2 = (20, 6) if G > 100 else (6, 3)
    cmc, ap = shard.rerank_eval_sharded(*a, k1, k2, 0.3, device=dev)
    merged = torch.from_numpy(np.concatenate([m["gf"], m["qf"]], 0)).to(dev)
    lab = [torch.from_numpy(m[k].astype(np.int32)).to(dev) for k in ("ql", "qc", "gl", "gc")]
    cmc1, ap1 = ops.rerank_eval(merged, G, *lab, k1, k2, 0.3)
    d_cmc, d_ap = float((cmc - cmc1).abs().max()), abs(float(ap) - float(ap1[0]))
    # (pair values outside a row's top-k1 list are direct dot products here and matrix entries on one GPU: 1e-7 apart,
    #  so a near-tie may swap two neighbours of one query)
    same = d_cmc <= 2.0 / Q and d_ap <= 2e-5
    ok &= same
    if rank == 0:
        print("shard_check world=%d rerank_eval Q=%d G=%d: row-sharded vs single max|dCMC| %.3g |dmAP| %.3g %s (rank1 %.6f mAP %.6f)" %
              (world, Q, G, d_cmc, d_ap, same, float(cmc[0]), float(ap)), flush=True)
    if G <= 2100:
        J = shard.compute_jaccard_distance_sharded(merged, k1, k2)
        J1 = ops.jaccard(merged, k1, k2)
        dj = float((J - J1).abs().max())
        block, (lo, hi) = shard.compute_jaccard_distance_sharded(merged, k1, k2, gather=False)
        same = dj <= 2e-6 and bool(torch.equal(block, J[lo:hi]))
        ok &= same
        Jn = faiss_utils.compute_jaccard_distance(merged, k1, k2, print_flag=False, search_option=2)  # the reference-named entry
        ok &= bool(np.array_equal(Jn, J.cpu().numpy()))
        if rank == 0:
            print("shard_check world=%d jaccard N=%d: row-sharded vs single max|dJ| %.3g %s" % (world, Q + G, dj, same), flush=True)
t = torch.tensor([0 if ok else 1], device=dev)
if world > 1:
    dist.all_reduce(t)
    dist.destroy_process_group()
sys.exit(int(t.item() != 0))
