"""MOT20-shaped Tracker._match with the device assignment: per-kernel CUDA-event times (for ncu captures and timings)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import reidb200
from reidb200 import _lib as L, matching, synth, tracker

M, T = 200, 300
fr = synth.mot_frame(M=M, T=T, budget=100, D=512, seed=7)
metric = matching.NearestNeighborDistanceMetric("cosine", 0.2, 100, max_tracks=T)
counts = [len(s) for s in fr["samples"]]
metric.partial_fit(np.concatenate(fr["samples"], 0), np.repeat(np.arange(T), counts), list(range(T)))
conf = np.arange(T) % 7 != 0
args = (metric, fr["det_feat"], fr["det_tlwh"], np.arange(T), conf, fr["trk_mean"], fr["trk_SI"], fr["trk_tsu"], 0.995, 0.7, 0.2)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 50
for _ in range(3):
    res = tracker.match_arrays(*args, device_lap=True)
L.profile_enable(True)
for _ in range(n):
    res = tracker.match_arrays(*args, device_lap=True)
torch.cuda.synchronize()
print({k: round(v[1] / v[0] * 1e3, 2) for k, v in L.profile_read().items()}, "us per launch")
L.profile_enable(False)
print("matches", len(res[0]))
