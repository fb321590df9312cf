"""Keep a rank's host buffers next to its GPU.

One process per GPU; every rank stages its shard in pinned host memory.  On a two-socket host a process
that runs on the far socket gets its pinned pages there (first touch) and every byte of its copies then
crosses the socket interconnect, which 8 ranks copying at once saturate long before PCIe.  Binding the
process to the CPUs the driver reports as local to its GPU, BEFORE the pinned buffers are allocated, keeps
the pages local."""
from __future__ import annotations

import os
from typing import List, Optional


def gpu_local_cpus(device_index: int) -> Optional[List[int]]:
    """CPUs NVML reports as local to the GPU (None when NVML or the information is unavailable)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        # CUDA_VISIBLE_DEVICES renumbers CUDA devices, NVML does not: go through the UUID-free path only
        # when no mask is set, else map the visible entry back to a physical index when it is numeric
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        phys = device_index
        if vis:
            parts = [p.strip() for p in vis.split(",") if p.strip()]
            if device_index < len(parts) and parts[device_index].isdigit():
                phys = int(parts[device_index])
            elif device_index < len(parts):
                h = pynvml.nvmlDeviceGetHandleByUUID(parts[device_index].encode()
                                                     if isinstance(parts[device_index], str) else parts[device_index])
                phys = pynvml.nvmlDeviceGetIndex(h)
        h = pynvml.nvmlDeviceGetHandleByIndex(phys)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = [64 * w + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1]
        return cpus or None
    except Exception:
        return None


def bind_to_gpu(device_index: int) -> Optional[List[int]]:
    """Restrict this process to the CPUs local to the GPU.  Returns the CPU list, or None if nothing changed
    (no NVML, no affinity information, or the local CPUs are not in this process's allowed set)."""
    cpus = gpu_local_cpus(device_index)
    if not cpus:
        return None
    try:
        allowed = os.sched_getaffinity(0)
        use = sorted(set(cpus) & allowed)
        if not use:
            return None
        os.sched_setaffinity(0, use)
        return use
    except Exception:
        return None
