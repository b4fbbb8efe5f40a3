"""Diagnostics: candidate-list statistics of the fused selection for a retrieval-shaped k-NN (REIDB200_SEL_STATS)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["REIDB200_SEL_STATS"] = "1"
os.environ["REIDB200_KNN_SELECT"] = "1"
import torch
import reidb200
from reidb200 import ops
nq, nb, d, k = [int(x) for x in (sys.argv[1:5] if len(sys.argv) > 4 else (10000, 1000000, 256, 100))]
g = torch.Generator(device="cuda").manual_seed(0)
xb = torch.nn.functional.normalize(torch.randn((nb, d), generator=g, device="cuda"), dim=1)
xq = torch.nn.functional.normalize(torch.randn((nq, d), generator=g, device="cuda"), dim=1)
for it in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    D, I = ops.knn(xq, xb, k, "sql2")
    torch.cuda.synchronize()
    print("knn %dx%dx%d top-%d: %.1f ms" % (nq, nb, d, k, (time.perf_counter() - t0) * 1e3))
