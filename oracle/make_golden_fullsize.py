"""BASELINE.json configs[0] / configs[1] at FULL size through the REFERENCE's own code -> tests/golden/market_full.npz.

TEST INFRASTRUCTURE ONLY; runs only in the build container (needs /root/reference), takes several minutes and a few
GB of host memory:

    python -m oracle.make_golden_fullsize [--noise 2.5] [--seed 3]

Input: synth.market_like(Q=3368, G=15913, D=512, n_ids=751, seed, noise) -- the generator is seeded numpy, so the GPU
box regenerates the same features (the fixture stores a SHA-256 of them and the test checks it).  noise = 2.5 is the
HARD setting (plain rank-1 ~ 0.95 / mAP ~ 0.63): first hits spread over many ranks, CMC is a curve.

What is stored (all produced by reference code imported in place):
  plain_*     reid/evaluate.py evaluate_all + per-query evaluate(): CMC, mAP, AP and first-hit rank of every query (C1)
  J rows      reid/faiss_utils.py compute_jaccard_distance(merged, k1=20, k2=6, search_option=3) on merged = cat(gallery,
              query) (image_reid_inference.py:270), through the exact-L2 faiss stand-in: sampled rows of the N x N matrix
  rr_*        the re-ranked evaluation: dist = (1 - lambda) * J[G:, :G] + lambda * euclidean_dist(qf, gf)
              (reid/losses/utils.py:21-35; the blend is the north_star addition), every row ranked by a
              stable argsort and reduced by the reference's compute_mAP (reid/evaluate.py:78-105) with the reference's
              good / junk rule (:66-72): CMC, mAP, per-query AP, first hit, and per query the smallest gap between a good
              item and a neighbouring non-good item in the ranking ("margin": below the distance tolerance a swap is a tie)
  rank rows   the initial ranking (exact fp32 squared-L2 k-NN stand-in, k = k1) of sampled rows with their distances
The oracle restatement (oracle/reid_oracle.py) is asserted equal on the same inputs before anything is written.
"""
from __future__ import annotations

import argparse
import contextlib
import hashlib
import io
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "real-time-reid-tracking_b200"))

from oracle import ref_import, reid_oracle as O  # noqa: E402
from oracle.make_golden import ranked_eval  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
Q, G, D, IDS = 3368, 15913, 512, 751


def feature_digest(d):
    h = hashlib.sha256()
    for k in ("qf", "ql", "qc", "gf", "gl", "gc"):
        h.update(np.ascontiguousarray(d[k]).tobytes())
    return h.hexdigest()


def main():
    ap_ = argparse.ArgumentParser()
    ap_.add_argument("--noise", type=float, default=2.5)
    ap_.add_argument("--seed", type=int, default=3)
    ap_.add_argument("--out", default="market_full")
    a = ap_.parse_args()
    assert ref_import.available(), "/root/reference not present"
    ref_import.install()
    import synth

    E = ref_import.load_file("ref_evaluate", "reid/evaluate.py")
    F = ref_import.load_file("ref_faiss_utils", "reid/faiss_utils.py")
    U = ref_import.load_file("ref_losses_utils", "reid/losses/utils.py")
    k1, k2, lam = 20, 6, 0.3
    d = synth.market_like(Q=Q, G=G, D=D, n_ids=IDS, seed=a.seed, noise=a.noise)
    t = {k: torch.from_numpy(v) for k, v in d.items()}
    out = dict(seed=np.int64(a.seed), noise=np.float64(a.noise), digest=np.array(feature_digest(d)), k1=np.int32(k1), k2=np.int32(k2),
               lambda_value=np.float64(lam))

    # ---- C1: the reference's evaluate_all and per-query evaluate
    t0 = time.time()
    with contextlib.redirect_stdout(io.StringIO()):
        cmc, mAP = E.evaluate_all(t["qf"], t["ql"], t["qc"], t["gf"], t["gl"], t["gc"])
    aps, first = np.zeros(Q), np.full(Q, -1, np.int32)
    for i in range(Q):
        ap_i, c = E.evaluate(t["qf"][i], t["ql"][i], t["qc"][i], t["gf"], t["gl"], t["gc"])
        if c[0] != -1:
            aps[i], first[i] = float(ap_i), int(np.argmax(c.numpy()))
    print("plain evaluate_all %.1f s: rank-1 %.6f mAP %.6f, first hits up to rank %d" % (time.time() - t0, cmc[0], mAP, first.max()))
    out.update(plain_cmc=cmc.numpy(), plain_map=np.float64(mAP), plain_ap_q=aps, plain_first=first)

    # ---- C2: reference compute_jaccard_distance on the merged set
    merged = torch.cat([t["gf"], t["qf"]], 0)
    t0 = time.time()
    J = F.compute_jaccard_distance(merged, k1=k1, k2=k2, print_flag=False, search_option=3)
    t_j = time.time() - t0
    print("reference compute_jaccard_distance(N=%d) %.1f s" % (merged.shape[0], t_j))
    mnp = merged.numpy()
    knn_d, rank = O.knn_l2(mnp, mnp, k1)
    orig = U.euclidean_dist(t["qf"], t["gf"]).numpy()
    dist = (J[G:, :G] * np.float32(1.0 - lam) + orig * np.float32(lam)).astype(np.float32)
    t0 = time.time()
    rr_cmc, rr_map, rr_ap, rr_first, rr_margin = ranked_eval(E, dist, d["ql"], d["qc"], d["gl"], d["gc"])
    print("re-ranked evaluation %.1f s: rank-1 %.6f mAP %.6f" % (time.time() - t0, rr_cmc[0], rr_map))
    # the oracle restatement on the same inputs (pins oracle/reid_oracle.py at full size)
    t0 = time.time()
    Jo = O.compute_jaccard_distance(mnp, k1, k2, initial_rank=rank)
    print("oracle compute_jaccard_distance %.1f s, max |J_oracle - J_reference| = %.3g" % (time.time() - t0, np.abs(Jo - J).max()))
    assert np.abs(Jo - J).max() <= 2e-6
    ocmc, oap = O.evaluate_all_from_dist(dist, d["ql"], d["qc"], d["gl"], d["gc"])
    assert np.array_equal(ocmc, rr_cmc) and abs(oap - rr_map) < 1e-12
    del Jo
    rng = np.random.default_rng(99)
    rows_q = np.sort(rng.choice(Q, 24, replace=False))
    rows_n = np.sort(rng.choice(G + Q, 8, replace=False))
    rows_r = np.sort(rng.choice(G + Q, 512, replace=False))
    out.update(jaccard_seconds=np.float64(t_j), rr_cmc=rr_cmc, rr_map=np.float64(rr_map), rr_ap_q=rr_ap, rr_first=rr_first,
               rr_margin=rr_margin, dist_rows=rows_q.astype(np.int32), dist_vals=dist[rows_q],
               J_rows=rows_n.astype(np.int32), J_vals=J[rows_n], rank_rows=rows_r.astype(np.int32),
               rank_vals=rank[rows_r].astype(np.int32), rank_dist=knn_d[rows_r].astype(np.float32),
               J_row_sums=J.sum(1, dtype=np.float64), J_nnz_lt1=np.int64((J < 1).sum()))
    os.makedirs(GOLD, exist_ok=True)
    path = os.path.join(GOLD, a.out + ".npz")
    np.savez_compressed(path, **out)
    print("wrote %s %.1f KB" % (path, os.path.getsize(path) / 1024))


if __name__ == "__main__":
    main()
