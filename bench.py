"""bench.py -- the reference's headline metric on B200: Q x G pairs/s of the Market1501-shaped evaluation INCLUDING
k-reciprocal re-ranking (BASELINE.json configs[1]: 3368 query x 15913 gallery, 512-d, k1=20 k2=6 lambda=0.3), plus the
MOT16/MOT20-shaped crop + association microseconds per frame (reported under "assoc" in the same JSON line).

  python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torch.distributed.run, one rank per GPU)
  python bench.py --impl reference --steps K --warmup W    (the reference algorithm's CPU port, oracle/, on the host cores)

A step = one pass of the hot path over one batch: features resident in HBM -> pack -> N x N squared-L2 tcgen05 GEMM ->
top-k1 -> k-reciprocal sets -> sparse V -> query expansion -> inverted index -> Jaccard + lambda blend (Q x G block) ->
CMC/mAP.  At N GPUs every rank evaluates its own Market-shaped set (independent evaluation sets: no data-path
collective, weak scaling); the timed region is bracketed by barrier + synchronize and the maximum over ranks is used.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

Q_FULL, G_FULL, D_FULL, IDS_FULL = 3368, 15913, 512, 751
K1, K2, LAMBDA = 20, 6, 0.3
METRIC = "QxG pairs/s incl. k-reciprocal re-rank (Market1501-shaped eval)"


def _peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.FIELDS,
                                          "--format=csv,noheader,nounits", "-lms", "50"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.06)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            p = [x.strip() for x in r.split(",")]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0])); mx.append(float(p[1]))
            except ValueError:
                continue
            for n, v in zip(names, p[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------------------------- our arm
def _dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def _timed(fn, steps, sync):
    """CUDA-event time (ms) of `steps` calls of fn on the current stream, synchronised on both sides."""
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    sync()
    return ms


def bench_assoc(steps, warmup):
    """MOT16 / MOT20-shaped crop + association, microseconds per frame (device resident; and end to end)."""
    import reidb200  # noqa: F401
    from reidb200 import _lib as L, matching, ops, preprocess, synth, tracker

    out = {}
    dev = torch.device("cuda", torch.cuda.current_device())
    for name, M, T in (("mot16", 30, 100), ("mot20", 200, 300)):
        fr = synth.mot_frame(M=M, T=T, budget=100, D=512, seed=7)
        H, W = fr["img"].shape[:2]
        xyxy = preprocess.xywh_to_xyxy(fr["det_xywh"], W, H)
        ok = (xyxy[:, 2] > xyxy[:, 0]) & (xyxy[:, 3] > xyxy[:, 1])
        xyxy = xyxy[ok]
        metric = matching.NearestNeighborDistanceMetric("cosine", 0.2, 100, max_tracks=T)
        counts = [len(s) for s in fr["samples"]]
        metric.partial_fit(np.concatenate(fr["samples"], 0), np.repeat(np.arange(T), counts), list(range(T)))
        frame = torch.from_numpy(fr["img"]).to(dev)
        off, geom = preprocess._geometry_from_xyxy(xyxy, W)
        off_d, geom_d = torch.from_numpy(off).to(dev), torch.from_numpy(geom).to(dev)
        crops = torch.empty((len(xyxy), 3, 256, 128), dtype=torch.float32, device=dev)
        det_feat = torch.from_numpy(fr["det_feat"]).to(dev)
        det_tlwh = torch.from_numpy(fr["det_tlwh"]).to(dev)
        slots = torch.from_numpy(metric.slots_for(np.arange(T))).to(dev)
        mean = torch.from_numpy(fr["trk_mean"]).to(dev)
        SI = torch.from_numpy(fr["trk_SI"].reshape(T, 16)).to(dev)
        tsu = torch.from_numpy(fr["trk_tsu"]).to(dev)
        app = torch.empty((T, M), dtype=torch.float64, device=dev)
        iou = torch.empty((T, M), dtype=torch.float64, device=dev)
        lib = L.load()
        mean3 = ops._f3((0.5, 0.5, 0.5))

        def frame_step():
            L.check(lib.reidb200_crop_resize_norm(frame.data_ptr(), off_d.data_ptr(), geom_d.data_ptr(), len(xyxy), 256, 128,
                                                  mean3, mean3, 0, crops.data_ptr(), L.stream_ptr()))
            L.check(lib.reidb200_assoc_frame(metric._gallery._h, det_feat.data_ptr(), det_feat.stride(0), det_tlwh.data_ptr(), M,
                                             slots.data_ptr(), mean.data_ptr(), SI.data_ptr(), tsu.data_ptr(), T, 0.995, 0.2, 0.7,
                                             0, 3, app.data_ptr(), None, iou.data_ptr(), L.stream_ptr()))

        n = max(steps * 20, 100)
        for _ in range(max(warmup, 3)):
            frame_step()
        L.profile_enable(True)
        ms = _timed(frame_step, n, torch.cuda.synchronize)
        prof = L.profile_read()
        L.profile_enable(False)
        # CUDA-graph replay of the same two launches (what a per-stream tracker loop would do)
        g = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            frame_step()
            torch.cuda.synchronize()
            with torch.cuda.graph(g, stream=s):
                frame_step()
        torch.cuda.synchronize()
        ms_graph = _timed(g.replay, n, torch.cuda.synchronize)
        # end to end through the public API: host boxes / detections / track state in, cost matrices out (host numpy).
        # tracker.FrameContext = two C calls per frame (one pinned H2D each, one D2H of both matrices, no Python marshalling)
        fc = tracker.FrameContext(metric, max_dets=256, max_tracks=512)
        xyxy32 = np.ascontiguousarray(xyxy, dtype=np.int32)
        ids, conf_all = np.arange(T), np.ones(T, bool)
        slots_h = metric.slots_for(ids)
        SI16 = np.ascontiguousarray(fr["trk_SI"].reshape(T, 16))

        def e2e_step():
            fc.crops(frame, xyxy32, out=crops)
            return fc.costs(det_feat, fr["det_tlwh"], ids, conf_all, fr["trk_mean"], SI16, fr["trk_tsu"], 0.995, 0.2, 0.7, slots=slots_h)

        def e2e_step_py():  # round 1's path: per-frame Python marshalling (tracker.frame_costs + costs_to_host)
            c = preprocess.preprocess_frame(frame, xyxy, out=crops)
            a, _, i = tracker.frame_costs(metric, det_feat, fr["det_tlwh"], np.arange(T), np.ones(T, bool), fr["trk_mean"],
                                          fr["trk_SI"], fr["trk_tsu"], 0.995, 0.2, 0.7)
            return tracker.costs_to_host(a, i)  # one D2H of both cost matrices, as tracker.match_arrays does

        a_fc, i_fc = e2e_step()
        a_py, i_py = e2e_step_py()
        assert np.array_equal(a_fc, a_py) and np.array_equal(i_fc, i_py), "FrameContext and frame_costs differ"
        e2e = {}
        for tag, fn in (("ctx", e2e_step), ("py", e2e_step_py)):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(n):
                fn()
            torch.cuda.synchronize()
            e2e[tag] = (time.perf_counter() - t0) / n * 1e6
        e2e_us = e2e["ctx"]
        # the whole Tracker._match from host arrays (costs + both assignment stages): Hungarian on the host (scipy, the
        # reference's solver) against the device assignment (N1)
        conf = np.arange(T) % 7 != 0
        margs = (metric, det_feat, fr["det_tlwh"], np.arange(T), conf, fr["trk_mean"], fr["trk_SI"], fr["trk_tsu"], 0.995, 0.7, 0.2)
        match_us = {}
        # device_lap = exact mode (scipy's own assignment, ties and list orders included: one warp runs _lsap.cpp on the whole
        # clamped matrix); device_lap_fast = the component-parallel solver (same matches, unmatched detections ascending)
        for tag, dl in (("host_lap", False), ("device_lap", True), ("device_lap_fast", "fast")):
            for _ in range(3):
                res = tracker.match_arrays(*margs, device_lap=dl)
            t0 = time.perf_counter()
            for _ in range(n):
                res = tracker.match_arrays(*margs, device_lap=dl)
            match_us[tag] = (time.perf_counter() - t0) / n * 1e6
            match_us[tag + "_matches"] = len(res[0])
        for dl, key, names in ((True, "lap_exact_kernels_us", ("lap_exact_kernel",)),
                               ("fast", "lap_kernels_us", ("lap_components_kernel", "lap_solve_kernel"))):
            L.profile_enable(True)
            for _ in range(20):
                tracker.match_arrays(*margs, device_lap=dl)
            torch.cuda.synchronize()
            lp = L.profile_read()
            L.profile_enable(False)
            match_us[key] = sum(lp[k][1] for k in names if k in lp) / 20 * 1e3
        crop_us = prof["crop_cv_linear_kernel"][1] / prof["crop_cv_linear_kernel"][0] * 1e3
        assoc_us = prof["assoc_frame_kernel"][1] / prof["assoc_frame_kernel"][0] * 1e3
        crop_bytes = len(xyxy) * 3 * 256 * 128 * 4 + int((geom[:, 0].astype(np.int64) * geom[:, 1] * 3).sum())
        gal_bytes = sum(counts) * 512 * 4 + M * 512 * 4 + 2 * T * M * 8
        out[name] = {"dets": int(M), "tracks": int(T), "us_per_frame": ms / n * 1e3, "us_per_frame_cuda_graph": ms_graph / n * 1e3,
                     "e2e_us_per_frame": e2e_us, "e2e_us_per_frame_python_marshalling": e2e["py"],
                     "e2e_h2d_bytes": int(len(xyxy) * 20 + M * 16 + T * 200), "e2e_d2h_bytes": int(2 * T * M * 8),
                     "crop_kernel_us": crop_us, "assoc_kernel_us": assoc_us,
                     "crop_gbs": crop_bytes / crop_us / 1e3, "assoc_gbs": gal_bytes / assoc_us / 1e3,
                     "assoc_tflops": 2.0 * sum(counts) * M * 512 / assoc_us / 1e6, "match_e2e_us": match_us}
    return out


def bench_next_rows(steps):
    """N2 / N3 rows at MOT20 size (300 tracks x 200 detections): kernel microseconds from CUDA events on the
    launch stream, and the same call of the numpy oracle on the host beside it."""
    import reidb200  # noqa: F401
    from oracle import reid_oracle as O
    from reidb200 import _lib as L, ops
    from reidb200.kalman import BatchedStrongSortKF

    rng = np.random.default_rng(3)
    T, M = 300, 200
    xy, wh = rng.uniform(0, 1800, (T + M, 2)), rng.uniform(8, 260, (T + M, 2))
    bx = np.concatenate([xy, xy + wh], 1)
    a, b = bx[:T], bx[T:]
    dev = torch.device("cuda", torch.cuda.current_device())
    a_d, b_d = torch.from_numpy(a).to(dev), torch.from_numpy(b).to(dev)
    out = torch.empty((T, M), dtype=torch.float64, device=dev)
    lib = L.load()
    n = max(20 * steps, 100)
    res = {"tracks": T, "dets": M}

    def kernel_us(fn, name):
        for _ in range(3):
            fn()
        L.profile_enable(True)
        _timed(fn, n, torch.cuda.synchronize)
        prof = L.profile_read()
        L.profile_enable(False)
        return prof[name][1] / prof[name][0] * 1e3

    def host_us(fn, reps=5):
        fn()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        return (time.perf_counter() - t0) / reps * 1e6

    for mode in ("iou", "giou", "diou", "ciou"):
        res["box_pairs_%s_us" % mode] = kernel_us(lambda: L.check(lib.reidb200_box_pairs(
            a_d.data_ptr(), T, b_d.data_ptr(), M, 1, L.BOX_MODES[mode], out.data_ptr(), L.stream_ptr())), "box_pairs_kernel")
    res["box_pairs_ciou_numpy_us"] = host_us(lambda: O.box_pairs(a, b, "ciou"))
    det = np.concatenate([b, rng.uniform(0.3, 1, (M, 1))], 1)
    trk = np.concatenate([a, np.zeros((T, 1))], 1)
    vel = rng.normal(0, 1, (T, 2))
    prev = np.concatenate([a - 5.0, np.ones((T, 1))], 1)
    d_d, t_d, v_d, p_d = (torch.from_numpy(x).to(dev) for x in (det, trk, vel, prev))
    o1, o2 = torch.empty((M, T), dtype=torch.float64, device=dev), torch.empty((M, T), dtype=torch.float64, device=dev)
    res["ocsort_cost_us"] = kernel_us(lambda: L.check(lib.reidb200_ocsort_cost(
        d_d.data_ptr(), M, t_d.data_ptr(), T, v_d.data_ptr(), p_d.data_ptr(), 0.2, o1.data_ptr(), o2.data_ptr(), L.stream_ptr())),
        "ocsort_cost_kernel")
    res["ocsort_cost_numpy_us"] = host_us(lambda: O.ocsort_cost(det, trk, vel, prev, 0.2))
    z = np.stack([rng.uniform(0, 1900, T), rng.uniform(0, 1000, T), rng.uniform(0.2, 0.7, T), rng.uniform(40, 400, T)], 1)
    kf = BatchedStrongSortKF(max_tracks=T)
    slots = torch.arange(T, dtype=torch.int32, device=dev)
    z_d, c_d = torch.from_numpy(z).to(dev), torch.full((T,), 0.5, dtype=torch.float64, device=dev)
    kf.initiate(slots, z_d)
    res["kf_predict_us"] = kernel_us(lambda: kf.predict(slots), "kf_predict_kernel")
    kf.initiate(slots, z_d)
    kf.predict(slots)
    res["kf_update_us"] = kernel_us(lambda: kf.update(slots, z_d, c_d), "kf_update_kernel")
    st = [O.kf_initiate(z[t]) for t in range(T)]

    def host_kf():
        for t in range(T):
            x, P = O.kf_predict(st[t][0], st[t][1])
            O.kf_update(x, P, z[t], 0.5)

    res["kf_predict_update_numpy_us"] = host_us(host_kf, 2)
    return res


def bench_posttricks(merged, k1, k2, steps):
    """N4 row at the Market-shaped merged set: DBSCAN / clamp + add on the device-resident N x N Jaccard matrix
    (ms, CUDA events on the launch stream) with sklearn's DBSCAN on a bounded host sample beside it."""
    import reidb200  # noqa: F401
    from reidb200 import _lib as L, ops

    N = merged.shape[0]
    J = ops.jaccard(merged, k1, k2)
    res = {"N": int(N), "matrix_bytes": int(N) * int(N) * 4}
    n = max(steps, 5)
    for _ in range(2):
        labels, ncl, core = ops.dbscan_precomputed(J, 0.5, 7)
    L.profile_enable(True)
    ms = _timed(lambda: ops.dbscan_precomputed(J, 0.5, 7), n, torch.cuda.synchronize)
    prof = L.profile_read()
    L.profile_enable(False)
    res["dbscan_ms"] = ms / n
    res["dbscan_kernel_ms"] = {k: v[1] / v[0] for k, v in prof.items()}
    deg = prof.get("dbscan_degree_kernel")
    if deg:
        res["dbscan_degree_read_GBps"] = res["matrix_bytes"] / (deg[1] / deg[0] * 1e-3) / 1e9
    res["dbscan_clusters"], res["dbscan_core"], res["dbscan_noise"] = int(ncl.item()), int(core.sum().item()), int((labels == -1).sum().item())
    other = torch.rand((N, N), dtype=torch.float32, device=J.device)
    res["clamp_add_ms"] = _timed(lambda: ops.clamp_add_(J, other), n, torch.cuda.synchronize) / n
    res["clamp_add_GBps"] = 3 * res["matrix_bytes"] / (res["clamp_add_ms"] * 1e-3) / 1e9
    try:
        from sklearn.cluster import DBSCAN
        S = 4000
        Jh = ops.jaccard(merged[:S].contiguous(), k1, k2).cpu().numpy()
        t0 = time.perf_counter()
        DBSCAN(eps=0.5, min_samples=7, metric="precomputed", n_jobs=-1).fit_predict(Jh)
        res["sklearn_dbscan_ms_sample"] = (time.perf_counter() - t0) * 1e3
        res["sklearn_sample"] = "%d x %d sub-matrix on the host cores (the reference's call, n_jobs=-1)" % (S, S)
    except ImportError:
        res["sklearn_dbscan_ms_sample"] = None
    return res


RETR_Q, RETR_G, RETR_D, RETR_K, RETR_PIECE = 10_000, 4_000_000, 2048, 100, 62_500


def _gallery_piece(piece, rows, d, dev):
    """Rows [piece*RETR_PIECE, ...) of the synthetic gallery: unit-norm N(0,I) rows, Philox-seeded by the GLOBAL piece id so
    that the gallery (and therefore the result table) is the same at every GPU count."""
    g = torch.Generator(device=dev)
    g.manual_seed(1_000_003 + piece)
    x = torch.randn((rows, d), generator=g, device=dev, dtype=torch.float32)
    return torch.nn.functional.normalize(x, dim=1)


def bench_retrieval(args, rank, world, dev, barrier):
    """BASELINE.json configs[4]: 10k queries x 4M gallery x 2048-d, top-100, gallery sharded over the ranks (strong scaling):
    resident packed shard per GPU -> tcgen05 GEMM + row-select per 2 GiB distance block -> one NCCL all-gather of the
    per-shard (distance, id)[Q,100] tables -> merge.  Device-timed (events, max over ranks) and end to end (pinned host
    queries in, host top-100 table out)."""
    import reidb200  # noqa: F401
    from reidb200 import _lib as L, shard

    G = args.retrieval_gallery
    Q, D, K = RETR_Q, RETR_D, RETR_K
    lo, hi = shard.shard_bounds(G, world)[rank]
    s = shard.ShardedSearcher(D, device=dev)
    s.reserve(hi - lo)
    r = lo
    while r < hi:  # piecewise: the fp32 rows of a piece are dropped as soon as they are packed
        piece, off = divmod(r, RETR_PIECE)
        rows = _gallery_piece(piece, RETR_PIECE, D, dev)[off:min(RETR_PIECE, off + hi - r)]
        s.append(rows)
        r += rows.shape[0]
    s.base = lo
    gq = torch.Generator().manual_seed(0)
    xq_host = torch.nn.functional.normalize(torch.randn((Q, D), generator=gq), dim=1).pin_memory()
    xq = xq_host.to(dev)
    steps = max(1, min(args.steps, 3))
    Dd, Id = s.search_device(xq, K)  # warm-up (NCCL communicator, kernel attributes)
    barrier()
    L.profile_enable(True)
    ms = _timed(lambda: s.search_device(xq, K), steps, barrier)
    prof = L.profile_read()
    L.profile_enable(False)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / steps
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        Dh, Ih = s.search(xq_host, K)
    t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t.item()) / steps
    gemm_n, gemm_ms = prof.get("gemm_split_tc_kernel", (0, 0.0))
    local_flops = 2.0 * Q * (hi - lo) * D * steps
    peaks, _ = _peaks()
    tf = local_flops / (gemm_ms * 1e-3) / 1e12 if gemm_ms else None
    return {"workload": "retrieval_%dq_x_%dg_d%d_top%d" % (Q, G, D, K), "scaling": "strong", "shards": world,
            "rows_per_shard": hi - lo, "ms_per_search": ms, "pairs_per_s": Q * G / (ms * 1e-3),
            "e2e_pairs_per_s": Q * G / e2e_s, "h2d_bytes_per_search": Q * D * 4, "d2h_bytes_per_search": Q * K * 12,
            "allgather_bytes_per_rank": Q * K * 12 if world > 1 else 0,
            "gemm_tflops_fp32_equiv": tf, "gemm_frac_of_peak_div3": (tf / (peaks["bf16_tflops_sustained"] / 3.0)) if tf else None,
            "kernel_ms": {k: round(v[1] / steps, 3) for k, v in sorted(prof.items(), key=lambda kv: -kv[1][1])},
            "launches": int(sum(n for n, _ in prof.values()) // steps),
            "checksum": {"id_sum": int(Ih.astype(np.int64).sum()), "dist_sum": float(Dh.astype(np.float64).sum())}}


def bench_rerank_sharded(args, rank, world, dev, barrier):
    """(e3) ONE re-ranked evaluation split over the ranks (shard.rerank_eval_sharded): k-NN rows sharded + all-gather of the
    ranking tables, sparse stages replicated, Jaccard + blend + CMC/mAP counting of this rank's query rows, one all-reduce.
    Strong scaling (the same set on every rank), device-timed, max over ranks.  Two workloads: configs[1] (hard data) --
    latency-bound at this size, reported for completeness -- and a merged set of 48000 rows (32000 queries x 16000 gallery rows: the fused
    evaluation keeps a query's gallery row in registers, <= 16384 columns), where the N x N work (9.2 GB of squared distances,
    never materialised as a whole) is what is being divided."""
    import reidb200  # noqa: F401
    from reidb200 import _lib as L, shard, synth

    out = {"scaling": "strong", "ranks": world,
           "collectives": "all_gather_into_tensor x2 (ranking tables i32 + f32 [N,k1]) + all_reduce(sum) f64[G+1]"}
    for tag, (Q, G, ids) in (("market", (Q_FULL, G_FULL, IDS_FULL)), ("large", (32000, 16000, 1900))):
        d = synth.market_like(Q=Q, G=G, D=D_FULL, n_ids=ids, seed=1, noise=HARD_NOISE)
        a = [torch.from_numpy(d[k]).to(dev) for k in ("qf", "ql", "qc", "gf", "gl", "gc")]
        fn = lambda: shard.rerank_eval_sharded(*a, K1, K2, LAMBDA, device=dev)
        for _ in range(2):
            cmc, ap = fn()
        barrier()
        steps = max(3, min(args.steps, 10))
        ms = _timed(fn, steps, barrier)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            import torch.distributed as dist
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item()) / steps
        # phases of one evaluation on this rank (CUDA events): k-NN rows, exchange of the tables, this rank's query rows
        merged = torch.cat([a[3], a[0]], 0)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        barrier()
        ev[0].record()
        ri, rd = shard.initial_rank_sharded(merged, K1)
        ev[1].record()
        cmc2, ap2 = fn()
        ev[2].record()
        torch.cuda.synchronize()
        N = Q + G
        row = {"queries": Q, "gallery": G, "N": N, "ms_per_eval": ms, "pairs_per_s": Q * G / (ms * 1e-3),
               "knn_stage_incl_allgather_ms": ev[0].elapsed_time(ev[1]),
               "allgather_bytes_per_rank": int(2 * ((N + world - 1) // world) * K1 * 4) if world > 1 else 0,
               "allreduce_bytes": (G + 1) * 8 if world > 1 else 0,
               "result": {"rank1": float(cmc[0]), "mAP": float(ap)}}
        out[tag] = row
        del a, d, merged
        torch.cuda.empty_cache()
    out["ms_per_eval"] = out["large"]["ms_per_eval"]
    out["note"] = ("market: N = 19281 is launch / latency bound (one GPU finishes the whole step in ~1.2 ms); large: N = 48000, "
                   "the k-NN stage (N^2 D flops / P per rank) dominates and divides")
    return out


def bench_eval_c1(args, rank, world, dev, barrier, d0):
    """BASELINE.json configs[0]: Market-shaped evaluate_all WITHOUT re-ranking (3368 x 15913 x 512-d score GEMM + CMC/mAP
    counting kernel), the queries split over the ranks and one NCCL all-reduce of the CMC counts + AP sum
    (shard.evaluate_all_sharded).  Strong scaling: the same set on every rank.  Device-timed, max over ranks."""
    from reidb200 import shard, synth

    d = d0 if d0 is not None else market_set(0, hard=True)
    a = [torch.from_numpy(d[k]).to(dev) for k in ("qf", "ql", "qc", "gf", "gl", "gc")]
    a[1], a[2], a[4], a[5] = (x.to(torch.int32) for x in (a[1], a[2], a[4], a[5]))
    fn = lambda: shard.evaluate_all_sharded(*a, device=dev)
    for _ in range(3):
        cmc, ap = fn()
    barrier()
    steps = max(args.steps, 5)
    ms = _timed(fn, steps, barrier)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / steps
    return {"workload": "market1501_eval_no_rerank (configs[0])", "scaling": "strong", "query_shards": world,
            "ms_per_eval": ms, "pairs_per_s": Q_FULL * G_FULL / (ms * 1e-3),
            "allreduce_bytes": (G_FULL + 1) * 8 if world > 1 else 0,
            "result": {"rank1": float(cmc[0]), "mAP": float(ap)}}


# --------------------------------------------------------------------------------------------- CPU port / reference arm
HARD_NOISE = 2.5   # Market-like difficulty (plain rank-1 ~ 0.87-0.95, mAP ~ 0.45-0.63); SURVEY 8d's recipe is noise = 0.6


def market_set(seed, hard=True):
    """BASELINE.json configs[0] / configs[1] input at FULL size: 3368 x 15913 x 512-d, 751 ids + junk, 6 cameras."""
    import reidb200  # noqa: F401
    from reidb200 import synth

    return synth.market_like(Q=Q_FULL, G=G_FULL, D=D_FULL, n_ids=IDS_FULL, seed=seed, noise=HARD_NOISE if hard else 0.6)


def cpu_step(d):
    """The reference algorithm on the host: exact k-NN, compute_jaccard_distance, lambda blend, CMC/mAP (oracle port)."""
    from oracle import reid_oracle as O

    dist = O.re_ranking_from_feats(d["qf"], d["gf"], K1, K2, LAMBDA)
    return O.evaluate_all_from_dist(dist, d["ql"], d["qc"], d["gl"], d["gc"])


def _threads(n):
    """Context manager limiting BLAS / OpenMP threads (the tracking scripts pin them to 1, track_yolov5.py:5-9)."""
    try:
        from threadpoolctl import threadpool_limits
        return threadpool_limits(limits=n)
    except Exception:
        import contextlib
        return contextlib.nullcontext()


def cpu_frame_us(M, T, threads, frames=12):
    """C3 / C4 on the host the way the reference runs a frame (sort/tracker.py:175-233): Extractor._preprocess of the M
    crops (cv2.resize when OpenCV is importable, else the oracle's restatement), NearestNeighborDistanceMetric.distance,
    gate_cost_matrix, iou_cost; the two Hungarian solves timed separately.  Median microseconds over `frames` frames."""
    from oracle import reid_oracle as O
    from reidb200 import preprocess, synth
    from scipy.optimize import linear_sum_assignment

    try:
        import cv2
        cv2.setNumThreads(threads if threads > 0 else -1)
    except Exception:
        cv2 = None
    fr = synth.mot_frame(M=M, T=T, budget=100, D=512, seed=7)
    H, W = fr["img"].shape[:2]
    xyxy = preprocess.xywh_to_xyxy(fr["det_xywh"], W, H)
    xyxy = xyxy[(xyxy[:, 2] > xyxy[:, 0]) & (xyxy[:, 3] > xyxy[:, 1])]
    crops = [fr["img"][b[1]:b[3], b[0]:b[2]] for b in xyxy]
    meas = O.tlwh_to_xyah_f32(fr["det_tlwh"])
    trk_tlwh = [O.track_to_tlwh(m[:4]) for m in fr["trk_mean"]]
    t_pre, t_cost, t_lap = [], [], []

    def pre():
        if cv2 is None:
            return O.preprocess_deepsort(crops)
        out = []
        for c in crops:  # feature_extractor.py:31-46
            im = cv2.resize(c.astype(np.float32) / 255., (128, 256))
            out.append(((im.transpose(2, 0, 1) - 0.5) / 0.5)[None])
        return np.concatenate(out, 0).astype(np.float32)

    with _threads(threads if threads > 0 else None):
        for f in range(frames + 1):
            t0 = time.perf_counter()
            pre()
            t1 = time.perf_counter()
            raw = O.nn_distance_matrix(fr["samples"], fr["det_feat"])
            app = O.gate_cost_matrix(raw, fr["trk_mean"][:, :4], fr["trk_SI"], meas, 0.995)
            iou = O.iou_cost(trk_tlwh, fr["trk_tsu"], fr["det_tlwh"])
            t2 = time.perf_counter()
            linear_sum_assignment(O.clamp_cost(app, 0.2))
            linear_sum_assignment(O.clamp_cost(iou, 0.7))
            t3 = time.perf_counter()
            if f:  # first frame = warm-up
                t_pre.append(t1 - t0); t_cost.append(t2 - t1); t_lap.append(t3 - t2)
    med = lambda v: float(np.median(v) * 1e6)
    return {"crop_us": med(t_pre), "cost_us": med(t_cost), "crop_plus_cost_us": med(np.add(t_pre, t_cost)), "hungarian_us": med(t_lap),
            "frames": frames, "crop_impl": "cv2.resize" if cv2 is not None else "numpy restatement"}


def cpu_rows(d_hard):
    """the host-side numbers that sit next to the GPU ones (numpy port of the reference, this box's
    cores): C1 evaluate_all, C3 / C4 per frame at 1 thread (as the tracking scripts pin) and all cores, C5 exact top-100
    over a 64 k sub-gallery extrapolated linearly to 4 M rows."""
    from oracle import reid_oracle as O

    cores = os.cpu_count() or 1
    out = {"cores": cores, "kind": "port"}
    keys = ("qf", "ql", "qc", "gf", "gl", "gc")
    for tag, n in (("1_thread", 1), ("all_cores", 0)):
        with _threads(n if n else None):
            t0 = time.perf_counter()
            cmc, ap = O.evaluate_all(*[d_hard[k] for k in keys])
            dt = time.perf_counter() - t0
        out.setdefault("c1_evaluate_all", {})[tag] = {"seconds": dt, "pairs_per_s": Q_FULL * G_FULL / dt}
        out["c1_evaluate_all"]["result"] = {"rank1": float(cmc[0]), "mAP": float(ap)}
        out.setdefault("c3_mot16_frame", {})[tag] = cpu_frame_us(30, 100, n)
        out.setdefault("c4_mot20_frame", {})[tag] = cpu_frame_us(200, 300, n, frames=8)
    rng = np.random.default_rng(0)
    sub = 65536
    xb = rng.standard_normal((sub, RETR_D), dtype=np.float32)
    xq = rng.standard_normal((RETR_Q, RETR_D), dtype=np.float32)
    t0 = time.perf_counter()
    O.knn_l2(xb, xq, RETR_K)
    dt = time.perf_counter() - t0
    out["c5_retrieval"] = {"sub_gallery": sub, "seconds_sub": dt, "seconds_4M_extrapolated": dt * RETR_G / sub,
                           "pairs_per_s": RETR_Q * sub / dt, "note": "10 k x 64 k x 2048 exact top-100 on all cores, extrapolated linearly in G"}
    return out


def run_reference(args):
    """The reference's own CPU algorithm for configs[1] at FULL size on the hard data (the numpy port: the Python reference
    cannot travel to the GPU box).  One step = one whole re-ranked evaluation (about a minute); the run stops after the
    requested steps or ~4 minutes, whichever comes first, and reports the steps it did."""
    rank, world, _ = _dist_env()
    if rank != 0:
        return
    torch.set_num_threads(os.cpu_count() or 1)
    d = market_set(0, hard=True)
    Q, G = Q_FULL, G_FULL
    t0 = time.perf_counter()
    steps = 0
    res = None
    for _ in range(max(args.steps, 1)):
        res = cpu_step(d)
        steps += 1
        if time.perf_counter() - t0 > 170:  # keep the whole run within a few minutes
            break
    dt = time.perf_counter() - t0
    v = steps * Q * G / dt
    sample = "FULL configs[1]: Market-shaped %dx%d, D=512, hard data (noise %.1f), numpy port of the reference algorithm" % (Q, G, HARD_NOISE)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "pairs/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": 0, "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": _config(1),
        "cpu_baseline": {"value": v, "unit": "pairs/s", "cores": os.cpu_count(), "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "result": {"rank1": float(res[0][0]), "mAP": float(res[1])},
        "note": "steps / warmup requested: %d / %d; a step is a whole full-size evaluation (~1 min), so the run is cut at ~3 min" % (args.steps, args.warmup),
    }))


def _config(world):
    return {"workload": "market1501_rerank_k20_6_l0.3_hard (configs[1]; noise %.1f: Market-like difficulty)" % HARD_NOISE,
            "queries": Q_FULL, "gallery": G_FULL, "dim": D_FULL, "k1": K1, "k2": K2, "lambda": LAMBDA,
            "parallelism": "independent evaluation sets per GPU" if world > 1 else "single GPU",
            "l2": "working set (N x N distance work + 0.5 GB sparse tables per step) larger than L2",
            "evaluation": "fused into the re-rank kernel (ops.rerank_eval); ms_per_step_materialised = jaccard + eval_cmc_map",
            "launch": "one CUDA graph replay per step (ops.RerankEvalPlan); ms_per_step_stream_launches = the same kernels launched one by one"}


def run_ours(args):
    import reidb200  # noqa: F401
    from reidb200 import _lib as L, evaluate, ops, synth

    rank, world, local = _dist_env()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        # NCCL prints its version banner on stdout when the communicator comes up: keep stdout to the one JSON line
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    L.require_device()
    # pinned staging buffers of the end-to-end path on the NUMA node of this rank's GPU (reidb200/numa.py); --no-numa-bind
    # leaves the process where the launcher put it
    from reidb200 import numa
    numa_info = {"node": None, "bound": False} if args.no_numa_bind else numa.bind_host_to_gpu(local)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    d = market_set(rank, hard=True)
    Q, G, N = Q_FULL, G_FULL, Q_FULL + G_FULL
    merged_host = torch.from_numpy(np.concatenate([d["gf"], d["qf"]], 0)).pin_memory()
    lab_host = {k: torch.from_numpy(d[k].astype(np.int32)).pin_memory() for k in ("ql", "qc", "gl", "gc")}
    merged = merged_host.to(dev)
    lab = {k: v.to(dev) for k, v in lab_host.items()}

    def step_launches():  # fused: the CTA that produces a query's re-ranked row reduces it to AP / first hit at once
        return ops.rerank_eval(merged, G, lab["ql"], lab["qc"], lab["gl"], lab["gc"], K1, K2, LAMBDA)

    # the same call captured once as a CUDA graph (ops.RerankEvalPlan): what evaluate.evaluate_all_reranked replays
    plan = ops.RerankEvalPlan(G, Q, D_FULL, K1, K2, LAMBDA, dev)
    plan.merged.copy_(merged)
    for dst, k in ((plan.ql, "ql"), (plan.qc, "qc"), (plan.gl, "gl"), (plan.gc, "gc")):
        dst.copy_(lab[k])
    step = plan.run

    def step_materialised():  # the same evaluation with the re-ranked Q x G matrix written and re-read (two calls)
        dist_qg = ops.jaccard(merged, K1, K2, False, LAMBDA, rows=(G, Q), cols=(0, G))
        return ops.eval_cmc_map(dist_qg, lab["ql"], lab["qc"], lab["gl"], lab["gc"], descending=False)

    for _ in range(max(args.warmup, 3)):
        cmc, ap = step()
    barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ms = _timed(step, args.steps, barrier)  # the headline: K steps, nothing but the product's launches on the stream
    # a second pass of K steps with every kernel bracketed by CUDA events on its launch stream (kernel shares, the
    # roofline kernel's launch time, the launch count); the event pairs cost a few per cent, hence the separate pass
    cmc_l, ap_l = step_launches()
    assert torch.equal(cmc_l, cmc) and torch.equal(ap_l, ap), "graph replay and plain launches differ"
    ms_launch = _timed(step_launches, args.steps, barrier)  # the same kernels launched one by one (no graph)
    L.profile_enable(True)
    ms_prof = _timed(step_launches, args.steps, barrier)
    prof = L.profile_read()
    L.profile_enable(False)
    clk = clocks.stop() if rank == 0 else None
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = world * args.steps * Q * G / (ms * 1e-3)
    cmc_m, ap_m = step_materialised()
    assert torch.equal(cmc_m, cmc) and torch.equal(ap_m, ap), "fused and materialised evaluations differ"
    ms_mat = _timed(step_materialised, args.steps, barrier) / args.steps

    # ---- end to end through the public API: pinned host features + labels in, CMC + mAP out, every step
    qf_h, gf_h = merged_host[G:], merged_host[:G]

    def e2e_step():
        return evaluate.evaluate_all_reranked(qf_h, lab_host["ql"], lab_host["qc"], gf_h, lab_host["gl"], lab_host["gc"], K1, K2,
                                              LAMBDA, verbose=False)

    for _ in range(2):
        cmc_h, ap_h = e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cmc_h, ap_h = e2e_step()
    torch.cuda.synchronize()
    t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_sync_value = world * args.steps * Q * G / float(t.item())

    # the same evaluations back to back through evaluate.RerankedEvalStream: every step still copies its inputs from
    # pinned host memory and reads its CMC + mAP back, but the copy of set i+1 overlaps the kernels of set i
    ev = evaluate.RerankedEvalStream(depth=2)

    def e2e_run(n):
        pending, out = None, None
        for _ in range(n):
            tk = ev.submit(qf_h, lab_host["ql"], lab_host["qc"], gf_h, lab_host["gl"], lab_host["gc"], K1, K2, LAMBDA)
            if pending is not None:
                out = ev.result(pending)
            pending = tk
        return ev.result(pending)

    cmc_p, ap_p = e2e_run(3)
    assert abs(ap_p - ap_h) < 1e-9 and torch.equal(cmc_p, cmc_h), "pipelined evaluation differs from the synchronous one"
    barrier()
    t0 = time.perf_counter()
    cmc_p, ap_p = e2e_run(args.steps)
    torch.cuda.synchronize()
    t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * args.steps * Q * G / float(t.item())
    h2d = merged_host.numel() * 4 + sum(v.numel() * 4 for v in lab_host.values())
    d2h = G * 4 + 8

    # ---- the same step on SURVEY 8d's recipe (noise 0.6: trivially separable, round 1's headline data) for continuity
    d_easy = market_set(rank, hard=False)
    plan.merged.copy_(torch.from_numpy(np.concatenate([d_easy["gf"], d_easy["qf"]], 0)))
    for dst, k in ((plan.ql, "ql"), (plan.qc, "qc"), (plan.gl, "gl"), (plan.gc, "gc")):
        dst.copy_(torch.from_numpy(d_easy[k].astype(np.int32)))
    for _ in range(3):
        cmc_e, ap_e = step()
    ms_easy = _timed(step, args.steps, barrier)
    t = torch.tensor([ms_easy], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_easy = float(t.item())
    easy = {"workload": "market1501_rerank_k20_6_l0.3 on SURVEY 8d's recipe (noise 0.6)", "ms_per_step": ms_easy / args.steps,
            "value": world * args.steps * Q * G / (ms_easy * 1e-3), "result": {"rank1": float(cmc_e[0]), "mAP": float(ap_e[0])}}
    del d_easy

    eval_c1 = bench_eval_c1(args, rank, world, dev, barrier, d if rank == 0 else None)
    rerank_sharded = bench_rerank_sharded(args, rank, world, dev, barrier) if not args.no_sharded_rerank else None
    posttricks = bench_posttricks(merged, K1, K2, args.steps) if (rank == 0 and not args.no_assoc) else None
    retrieval = None
    if not args.no_retrieval:
        merged = None
        plan = None
        torch.cuda.empty_cache()
        retrieval = bench_retrieval(args, rank, world, dev, barrier)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peaks, peak_kind = _peaks()
    gemm_n, gemm_ms = prof.get("gemm_split_tc_kernel", (0, 0.0))
    launches = sum(n for n, _ in prof.values())
    share = {k: round(v[1] / ms_prof, 4) for k, v in sorted(prof.items(), key=lambda kv: -kv[1][1])}
    # algorithmic (fp32-equivalent) flops of one launch: 2*D per UNORDERED pair -- the self-distance matrix is symmetric
    # and the kernel computes only its upper triangle of tiles
    gemm_flops = 2.0 * D_FULL * (N * (N + 1) / 2.0)
    try:  # dram__bytes_read.sum + dram__bytes_write.sum of this kernel from the committed ncu --set full capture
        traffic = json.load(open(os.path.join(ROOT, "profiles", "gemm_traffic.json")))["dram_bytes_per_launch"]
    except Exception:
        traffic = None
    achieved = gemm_flops / (gemm_ms / max(gemm_n, 1) * 1e-3) / 1e12 if gemm_n else None
    peak = peaks["bf16_tflops_sustained"] / 3.0  # three 16-bit UMMAs per fp32-accurate product
    line = {
        "metric": METRIC, "value": value, "unit": "pairs/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        # compact per-N numbers first (a long line may be cut at its tail): strong-scaling rows with a data-path collective
        "retrieval_ms_per_search": retrieval["ms_per_search"] if retrieval else None,
        "rerank_sharded_ms": rerank_sharded["ms_per_eval"] if rerank_sharded else None,
        "config": _config(world),
        "survey_recipe": easy,
        "ms_per_step_stream_launches": ms_launch / args.steps,
        "ms_per_step_materialised": ms_mat, "ms_per_step_with_kernel_events": ms_prof / args.steps,
        "clocks": clk,
        "e2e": {"value": e2e_value, "unit": "pairs/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "mode": "evaluate.RerankedEvalStream(depth=2): H2D of set i+1 overlaps the kernels of set i",
                "sync_call_value": e2e_sync_value, "host_numa": numa_info},
        "gpu_launches": int(launches),
        "roofline": {"kernel": "gemm_split_tc_kernel", "bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                     "frac": (achieved / peak) if achieved else None, "traffic": traffic,
                     "peak_source": "%s bf16_tflops_sustained / 3 (fp16 hi|lo split: three kind::f16 UMMAs per fp32-accurate product; fp16 and bf16 run at the same rate)" % peak_kind,
                     "units_per_launch": "N(N+1)/2 = %d unordered pairs x 2*D flop" % (N * (N + 1) // 2),
                     "tensor_pipe_tflops": achieved * 3 if achieved else None},
        "kernel_share_of_step": share,
        "result": {"rank1": float(cmc_h[0]), "mAP": float(ap_h)},
    }
    if rerank_sharded is not None:
        line["rerank_sharded"] = rerank_sharded
    line["eval_c1"] = eval_c1
    if retrieval is not None:
        line["retrieval"] = retrieval
    if not args.no_assoc:
        line["assoc"] = bench_assoc(args.steps, args.warmup)
        line["next_rows"] = bench_next_rows(args.steps)
        line["next_rows"]["n4"] = posttricks
    if not args.no_cpu and world == 1:
        t0 = time.perf_counter()
        res = cpu_step(d)  # FULL configs[1] on the same hard data, one pass (about a minute)
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": Q * G / dt, "unit": "pairs/s", "cores": os.cpu_count(), "kind": "port", "seconds": dt,
                                "sample": "FULL configs[1] (%dx%d, D=512, hard data), one pass, numpy port of the reference" % (Q, G),
                                "result": {"rank1": float(res[0][0]), "mAP": float(res[1])}}
        if not args.no_cpu_rows:
            line["cpu_rows"] = cpu_rows(d)
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-assoc", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-retrieval", action="store_true")
    ap.add_argument("--no-cpu-rows", action="store_true", help="skip the C1 / C3 / C4 / C5 host baselines (BASELINE.md section 2)")
    ap.add_argument("--no-sharded-rerank", action="store_true")
    ap.add_argument("--no-numa-bind", action="store_true", help="do not bind the process to its GPU's NUMA node")
    ap.add_argument("--retrieval-gallery", type=int, default=RETR_G)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
