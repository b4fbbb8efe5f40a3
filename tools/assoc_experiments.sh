#!/bin/bash
# timing-only experiments on the packed-tile association kernel (MOT20 shape): column-range width, stages, debug switches
run() {
  env "$@" timeout 300 python - <<'PY'
import os, sys
sys.path.insert(0, ".")
import numpy as np, torch
import reidb200
from reidb200 import _lib as L, matching, synth
M, T = 200, 300
fr = synth.mot_frame(M=M, T=T, budget=100, D=512, seed=7)
metric = matching.NearestNeighborDistanceMetric("cosine", 0.2, 100, max_tracks=T)
metric.partial_fit(np.concatenate(fr["samples"], 0), np.repeat(np.arange(T), [len(s) for s in fr["samples"]]), list(range(T)))
dev = torch.device("cuda")
f = lambda a, dt=None: torch.from_numpy(np.ascontiguousarray(a if dt is None else a.astype(dt))).to(dev)
det_feat, det_tlwh, slots = f(fr["det_feat"]), f(fr["det_tlwh"]), f(metric.slots_for(np.arange(T)))
mean, SI, tsu = f(fr["trk_mean"]), f(fr["trk_SI"].reshape(T, 16)), f(fr["trk_tsu"])
app = torch.empty((T, M), dtype=torch.float64, device=dev); iou = torch.empty_like(app)
lib = L.load()
def step():
    L.check(lib.reidb200_assoc_frame(metric._gallery._h, det_feat.data_ptr(), det_feat.stride(0), det_tlwh.data_ptr(), M, slots.data_ptr(),
                                     mean.data_ptr(), SI.data_ptr(), tsu.data_ptr(), T, 0.995, 0.2, 0.7, 0, 3, app.data_ptr(), None, iou.data_ptr(), L.stream_ptr()))
for _ in range(5): step()
L.profile_enable(True)
for _ in range(200): step()
torch.cuda.synchronize()
p = L.profile_read()
print({k: round(v[1] / v[0] * 1e3, 2) for k, v in p.items()})
PY
}
for cfg in "X=0" "REIDB200_ASSOC_BN=208" "REIDB200_ASSOC_BN=64" "REIDB200_ASSOC_BN=80" "REIDB200_ASSOC_STAGES=2" "REIDB200_ASSOC_DEBUG=1" "REIDB200_ASSOC_DEBUG=2" "REIDB200_ASSOC_DEBUG=3" "REIDB200_ASSOC_DEBUG=4" "REIDB200_ASSOC_DEBUG=7" "REIDB200_ASSOC_V1=1"; do
  echo -n "$cfg  "; run $cfg 2>&1 | tail -1
done
