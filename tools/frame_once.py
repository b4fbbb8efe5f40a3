"""A few MOT20-shaped frames of crop + association (for ncu captures of crop_cv_linear_kernel / assoc_frame_kernel)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import reidb200
from reidb200 import matching, preprocess, synth, tracker

M, T = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (200, 300)
fr = synth.mot_frame(M=M, T=T, budget=100, D=512, seed=7)
H, W = fr["img"].shape[:2]
xyxy = preprocess.xywh_to_xyxy(fr["det_xywh"], W, H)
xyxy = xyxy[(xyxy[:, 2] > xyxy[:, 0]) & (xyxy[:, 3] > xyxy[:, 1])]
metric = matching.NearestNeighborDistanceMetric("cosine", 0.2, 100, max_tracks=T)
counts = [len(s) for s in fr["samples"]]
metric.partial_fit(np.concatenate(fr["samples"], 0), np.repeat(np.arange(T), counts), list(range(T)))
frame = torch.from_numpy(fr["img"]).cuda()
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    crops = preprocess.preprocess_frame(frame, xyxy)
    a, _, i = tracker.frame_costs(metric, fr["det_feat"], fr["det_tlwh"], np.arange(T), np.ones(T, bool), fr["trk_mean"], fr["trk_SI"],
                                  fr["trk_tsu"], 0.995, 0.2, 0.7)
torch.cuda.synchronize()
print("ok", float(a.sum()), float(crops.sum()))
