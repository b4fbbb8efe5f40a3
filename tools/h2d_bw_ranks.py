"""torchrun --nproc-per-node N tools/h2d_bw_ranks.py : pinned host -> device bandwidth of N ranks copying AT THE SAME TIME
(the end-to-end evaluation's 40 MB per step), with the pinned buffer allocated (a) wherever the process happened to run and
(b) after binding the process to its GPU's NUMA node (reidb200.numa).  Prints one JSON line with per-rank and aggregate GB/s."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import reidb200  # noqa: F401
from reidb200 import numa

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def measure(tag):
    n = 40 * 1024 * 1024
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    h.fill_(1)
    d = torch.empty(n, dtype=torch.uint8, device=dev)
    for _ in range(3):
        d.copy_(h, non_blocking=True)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        d.copy_(h, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    gbs = n * 50 / (e0.elapsed_time(e1) * 1e-3) / 1e9
    t = torch.tensor([gbs], dtype=torch.float64, device=dev)
    allv = [torch.zeros_like(t) for _ in range(world)]
    if world > 1:
        dist.all_gather(allv, t)
    else:
        allv = [t]
    barrier()
    return {tag + "_per_rank_GBps": [round(float(x), 1) for x in allv], tag + "_aggregate_GBps": round(sum(float(x) for x in allv), 1)}


out = {"ranks": world, "cpus_before": len(os.sched_getaffinity(0))}
out.update(measure("unbound"))
info = numa.bind_host_to_gpu(local)
infos = [None] * world
if world > 1:
    dist.all_gather_object(infos, info)
else:
    infos = [info]
out["binding"] = infos
out.update(measure("numa_bound"))
if rank == 0:
    print(json.dumps(out), flush=True)
if world > 1:
    dist.destroy_process_group()
