"""Pinned host -> device copy bandwidth at the sizes the end-to-end evaluation moves (context for bench.py's e2e line)."""
import json, sys, torch
res = {}
for mb in (1, 8, 40, 160):
    n = mb * 1024 * 1024
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        d.copy_(h, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    res["h2d_%dMB_GBps" % mb] = n * 20 / (e0.elapsed_time(e1) * 1e-3) / 1e9
    e0.record()
    for _ in range(20):
        h.copy_(d, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    res["d2h_%dMB_GBps" % mb] = n * 20 / (e0.elapsed_time(e1) * 1e-3) / 1e9
print(json.dumps(res))
# the same 40 MB split over several streams (do concurrent copies add up?)
n = 40 * 1024 * 1024
h = torch.empty(n, dtype=torch.uint8).pin_memory()
h.fill_(1)
d = torch.empty(n, dtype=torch.uint8, device="cuda")
res2 = {}
for parts in (1, 2, 4):
    streams = [torch.cuda.Stream() for _ in range(parts)]
    sz = n // parts
    def go():
        for i, s in enumerate(streams):
            with torch.cuda.stream(s):
                d[i * sz:(i + 1) * sz].copy_(h[i * sz:(i + 1) * sz], non_blocking=True)
    for _ in range(3):
        go()
    torch.cuda.synchronize()
    import time
    t0 = time.perf_counter()
    for _ in range(20):
        go()
    torch.cuda.synchronize()
    res2["h2d_40MB_%d_streams_GBps" % parts] = n * 20 / (time.perf_counter() - t0) / 1e9
print(json.dumps(res2))
