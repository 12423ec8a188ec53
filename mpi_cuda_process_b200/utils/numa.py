"""Keep a rank's host side next to its GPU: CPU affinity (and with it the NUMA node of every later pinned allocation)
from the GPU's PCI locality.

Why it matters here: the end-to-end path uploads each step's slab from pinned host memory while the previous solve runs.
A launcher that does not bind its children (torchrun does not) leaves each rank on an arbitrary socket, and pinned pages
come from the node the allocating thread happens to run on (Linux' default local policy) -- on a two-socket 8-GPU box
about half of the ranks then feed their GPU across the socket interconnect.  Binding the process to the GPU's
``local_cpulist`` before the first pinned allocation puts buffers, the copy-submitting thread and the GPU on one node.

The reference has the opposite problem built in: every rank uses GPU 0 (no cudaSetDevice, MDF_kernel.cu:114-121) and
uploads the whole grid from pageable memory twice per generation (MDF_kernel.cu:161,171).

Nothing here is required for correctness; every step is best effort and reports what it did.
"""
from __future__ import annotations

import os
from typing import Dict, Optional, Set

SYSFS_PCI = "/sys/bus/pci/devices"


def parse_cpulist(text: str) -> Set[int]:
    """'0-3,8,10-11' -> {0,1,2,3,8,10,11} (the kernel's cpulist format)."""
    cpus: Set[int] = set()
    for part in text.strip().split(","):
        part = part.strip()
        if not part:
            continue
        if "-" in part:
            lo, hi = part.split("-", 1)
            cpus.update(range(int(lo), int(hi) + 1))
        else:
            cpus.add(int(part))
    return cpus


def pci_address(device_index: int) -> Optional[str]:
    """'dddd:bb:dd.0' of a CUDA device as torch sees it (honours CUDA_VISIBLE_DEVICES), or None."""
    try:
        import torch
        p = torch.cuda.get_device_properties(device_index)
        dom, bus, dev = getattr(p, "pci_domain_id", None), getattr(p, "pci_bus_id", None), getattr(p, "pci_device_id", None)
        if bus is not None and dev is not None:
            return f"{int(dom or 0):04x}:{int(bus):02x}:{int(dev):02x}.0"
    except Exception:
        pass
    try:                                              # NVML numbers devices without CUDA_VISIBLE_DEVICES
        if os.environ.get("CUDA_VISIBLE_DEVICES"):
            return None
        import pynvml
        pynvml.nvmlInit()
        bus_id = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(device_index)).busId
        bus_id = bus_id.decode() if isinstance(bus_id, bytes) else str(bus_id)
        dom, rest = bus_id.lower().split(":", 1)
        return f"{dom[-4:]}:{rest}"
    except Exception:
        return None


def local_cpus(pci: str, sysfs: str = SYSFS_PCI) -> Optional[Set[int]]:
    try:
        with open(os.path.join(sysfs, pci, "local_cpulist")) as f:
            cpus = parse_cpulist(f.read())
        return cpus or None
    except (OSError, ValueError):
        return None


def numa_node(pci: str, sysfs: str = SYSFS_PCI) -> Optional[int]:
    try:
        with open(os.path.join(sysfs, pci, "numa_node")) as f:
            return int(f.read().strip())
    except (OSError, ValueError):
        return None


def bind_to_gpu(device_index: int, sysfs: str = SYSFS_PCI, pci: Optional[str] = None) -> Dict:
    """Restrict this process to the CPUs next to CUDA device ``device_index``.  Returns a record of what happened:
    ``{"bound": bool, "pci": ..., "numa_node": ..., "cpus": n, "why": ...}``; never raises."""
    out: Dict = {"bound": False}
    try:
        pci = pci or pci_address(device_index)
        if pci is None:
            out["why"] = "no PCI address for the device"
            return out
        out["pci"] = pci
        out["numa_node"] = numa_node(pci, sysfs)
        cpus = local_cpus(pci, sysfs)
        if not cpus:
            out["why"] = "no local_cpulist in sysfs"
            return out
        allowed = os.sched_getaffinity(0)
        target = cpus & allowed
        if not target:
            out["why"] = "the GPU's CPUs are outside this process's cpuset"
            return out
        if target == allowed:
            out.update(bound=True, cpus=len(target), why="already local (one node, or bound by the launcher)")
            return out
        os.sched_setaffinity(0, target)
        out.update(bound=True, cpus=len(target))
    except Exception as e:          # affinity is an optimisation: report, never fail the run
        out["why"] = f"{type(e).__name__}: {e}"
    return out
