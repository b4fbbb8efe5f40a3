"""cProfile of the per-frame public API path (preprocess_frame + tracker.frame_costs + one D2H), MOT16 shape."""
import cProfile, os, pstats, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import reidb200
from reidb200 import matching, preprocess, synth, tracker
M, T = 30, 100
fr = synth.mot_frame(M=M, T=T, budget=100, D=512, seed=7)
H, W = fr["img"].shape[:2]
xyxy = preprocess.xywh_to_xyxy(fr["det_xywh"], W, H)
xyxy = xyxy[(xyxy[:, 2] > xyxy[:, 0]) & (xyxy[:, 3] > xyxy[:, 1])]
metric = matching.NearestNeighborDistanceMetric("cosine", 0.2, 100, max_tracks=T)
counts = [len(s) for s in fr["samples"]]
metric.partial_fit(np.concatenate(fr["samples"], 0), np.repeat(np.arange(T), counts), list(range(T)))
dev = torch.device("cuda", 0)
frame = torch.from_numpy(fr["img"]).to(dev)
crops = torch.empty((len(xyxy), 3, 256, 128), dtype=torch.float32, device=dev)
det_feat = torch.from_numpy(fr["det_feat"]).to(dev)
ids, conf = np.arange(T), np.ones(T, bool)
def step():
    preprocess.preprocess_frame(frame, xyxy, out=crops)
    a, _, i = tracker.frame_costs(metric, det_feat, fr["det_tlwh"], ids, conf, fr["trk_mean"], fr["trk_SI"], fr["trk_tsu"], 0.995, 0.2, 0.7)
    return tracker.costs_to_host(a, i)
for _ in range(20): step()
pr = cProfile.Profile(); pr.enable()
for _ in range(300): step()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
