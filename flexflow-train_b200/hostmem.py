"""Host-side staging helpers for the reference-facing path with HOST buffers.

The reference hands kernels device pointers only; whoever feeds them (its data loaders under
lib/runtime, out of scope here) stages batches through pinned memory. On a two-socket B200 host
the PCIe bandwidth of a pinned buffer depends on which NUMA node its pages landed on (measured
here: 31-55 GB/s for the same 64 MiB copy), so the staging buffers are first-touched from a CPU of
the GPU's own node."""
import os
from pathlib import Path
from typing import List, Optional

import torch


def _parse_cpulist(text: str) -> List[int]:
    cpus: List[int] = []
    for part in text.strip().split(","):
        if not part:
            continue
        if "-" in part:
            lo, hi = part.split("-")
            cpus.extend(range(int(lo), int(hi) + 1))
        else:
            cpus.append(int(part))
    return cpus


def gpu_numa_node(device_index: int) -> Optional[int]:
    """NUMA node of the GPU's PCIe function, or None when the platform does not expose one."""
    try:
        p = torch.cuda.get_device_properties(device_index)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        node = int(Path(f"/sys/bus/pci/devices/{bdf}/numa_node").read_text())
        return node if node >= 0 else None
    except Exception:
        return None


def bind_host_thread_to_gpu_node(device_index: int) -> dict:
    """Restricts the calling process to the CPUs of the GPU's NUMA node (first-touch then places
    pinned staging buffers on that node). Returns what was done, for the bench's config line."""
    node = gpu_numa_node(device_index)
    info = {"gpu_numa_node": node, "bound": False}
    if node is None:
        return info
    try:
        cpus = set(_parse_cpulist(Path(f"/sys/devices/system/node/node{node}/cpulist").read_text()))
        allowed = cpus & set(os.sched_getaffinity(0))
        if allowed:
            os.sched_setaffinity(0, allowed)
            info["bound"] = True
            info["cpus"] = len(allowed)
    except Exception:
        pass
    return info


def pinned_empty(shape, dtype) -> torch.Tensor:
    """Pinned host tensor whose pages are touched (and so placed) by the calling thread."""
    return torch.empty(shape, dtype=dtype, pin_memory=True)
