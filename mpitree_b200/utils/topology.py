"""Host topology helpers: run a rank on the CPUs next to its GPU.

Pinned host buffers are placed by first touch, and a host-to-device copy from the far socket crosses
the inter-socket link.  With eight ranks streaming their shards at once that link, not PCIe, becomes
the limit (41 ms instead of ~21 ms for the end-to-end fit of 100M x 64 at eight GPUs in round 2), so a
rank that feeds its GPU from host memory should bind itself *before* it allocates those buffers.
The reference has no counterpart (mpirun's binding options played that role for its CPU ranks).
"""
from __future__ import annotations

import os
from typing import List, Optional


def gpu_cpu_affinity(device_index: int) -> Optional[List[int]]:
    """CPUs NVML reports as local to GPU ``device_index`` (``CUDA_VISIBLE_DEVICES`` order), or ``None``."""
    try:
        import pynvml

        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        phys = device_index
        if vis:
            try:
                phys = int(vis.split(",")[device_index])
            except (ValueError, IndexError):
                phys = device_index
        handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
        n_cpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (n_cpu + 63) // 64)
        cpus = [64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1]
        return [c for c in cpus if c < n_cpu] or None
    except Exception:  # noqa: BLE001  (no NVML, no permission, container without the topology files)
        return None


def parse_cpulist(text: str) -> List[int]:
    """``"0-3,8,10-11"`` (the kernel's cpulist format) -> ``[0, 1, 2, 3, 8, 10, 11]``."""
    cpus: List[int] = []
    for part in text.strip().split(","):
        part = part.strip()
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.extend(range(int(lo), int(hi or lo) + 1))
    return cpus


def gpu_cpu_affinity_sysfs(device_index: int, sysfs_root: str = "/sys/bus/pci/devices") -> Optional[List[int]]:
    """The same from the kernel's PCI topology (``<device>/local_cpulist``) for the GPU torch calls
    ``device_index``: works where NVML will not report an affinity (some containers)."""
    try:
        import torch

        p = torch.cuda.get_device_properties(device_index)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        with open(os.path.join(sysfs_root, bdf, "local_cpulist")) as fh:
            cpus = parse_cpulist(fh.read())
        return cpus or None
    except Exception:  # noqa: BLE001  (no CUDA device, no sysfs entry, unreadable file)
        return None


def bind_to_gpu_numa(device_index: int) -> Optional[List[int]]:
    """Restrict this process to the CPUs local to its GPU; returns them (``None``: nothing was changed).

    Only CPUs the process is already allowed to use are kept, so a container's CPU set is respected.
    """
    cpus = gpu_cpu_affinity(device_index) or gpu_cpu_affinity_sysfs(device_index)
    if not cpus or not hasattr(os, "sched_setaffinity"):
        return None
    try:
        allowed = os.sched_getaffinity(0)
        keep = sorted(set(cpus) & set(allowed))
        if not keep or len(keep) == len(allowed):
            return None
        os.sched_setaffinity(0, keep)
        return keep
    except OSError:
        return None
