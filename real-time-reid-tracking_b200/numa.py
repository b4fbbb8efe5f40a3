"""Host side of a multi-GPU node: keep a rank's pinned staging buffers on the NUMA node its GPU hangs off.

One process per GPU moves ~40 MB of features per evaluation from pinned host memory; on a two-socket box with eight GPUs the
copies of the four GPUs behind the OTHER socket cross the inter-socket link unless the pages were allocated locally.  Linux
places pinned pages on the node of the allocating thread, so binding the process to the GPU's node BEFORE the first pinned
allocation is all it takes.  Nothing here touches the device; a box without the sysfs entries is left alone."""
from __future__ import annotations

import os


def _read(path):
    try:
        with open(path) as f:
            return f.read().strip()
    except OSError:
        return None


def _parse_cpulist(s):
    cpus = set()
    for part in (s or "").split(","):
        part = part.strip()
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            cpus.update(range(int(a), int(b) + 1))
        else:
            cpus.add(int(part))
    return cpus


def gpu_numa_node(index: int):
    """NUMA node of cuda:<index> (None when the platform does not say)."""
    import torch

    p = torch.cuda.get_device_properties(index)
    try:
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
    except AttributeError:
        return None
    v = _read("/sys/bus/pci/devices/%s/numa_node" % bdf)
    if v is None:
        return None
    node = int(v)
    return node if node >= 0 else None


def bind_host_to_gpu(index: int) -> dict:
    """Restrict this process to the CPUs of cuda:<index>'s NUMA node (so that pinned buffers allocated afterwards are local
    to the GPU).  Returns {"node": n or None, "cpus": count, "bound": bool}."""
    info = {"node": None, "cpus": len(os.sched_getaffinity(0)), "bound": False}
    try:
        node = gpu_numa_node(index)
    except Exception:
        return info
    if node is None:
        return info
    cpus = _parse_cpulist(_read("/sys/devices/system/node/node%d/cpulist" % node)) & os.sched_getaffinity(0)
    info["node"] = node
    if cpus:
        os.sched_setaffinity(0, cpus)
        info["cpus"] = len(cpus)
        info["bound"] = True
    return info
