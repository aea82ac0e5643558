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


def bind_to_gpu_numa(device_index: int) -> Optional[List[int]]:
    """Restrict this process to the CPUs local to its GPU; returns them (``None``: nothing was changed).

    Only CPUs the process is already allowed to use are kept, so a container's CPU set is respected.
    """
    cpus = gpu_cpu_affinity(device_index)
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
