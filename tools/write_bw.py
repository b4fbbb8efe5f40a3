"""Measures write-only, read-only and copy HBM bandwidth with torch library kernels (context for the GEMM's output-store bound)."""
import torch, json
x = torch.empty(1487 * 1024 * 1024 // 4, device="cuda", dtype=torch.float32)
y = torch.empty_like(x)
def t(fn, n=10):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
b = x.numel() * 4
r = {"bytes": b}
ms = t(lambda: x.fill_(1.0)); r["fill_GBs"] = b / ms / 1e6
ms = t(lambda: x.zero_()); r["memset_GBs"] = b / ms / 1e6
ms = t(lambda: y.copy_(x)); r["copy_GBs_rw"] = 2 * b / ms / 1e6
ms = t(lambda: x.sum()); r["read_GBs"] = b / ms / 1e6
print(json.dumps(r))
