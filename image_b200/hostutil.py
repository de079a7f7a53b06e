"""Host-side plumbing for multi-GPU runs (no compute)."""
from __future__ import annotations

import os


def bind_cpu_affinity(device_index: int) -> list[int] | None:
    """Pin the calling process to the CPUs NVML reports as local to GPU `device_index`
    (`nvmlDeviceGetCpuAffinity`), so that pinned host buffers allocated afterwards are first touched
    on that GPU's NUMA node.  With one process per GPU on a two-socket box this keeps every rank's
    H2D / D2H traffic off the inter-socket link.  Returns the CPU list, or None when NVML or the
    affinity call is unavailable (the run then proceeds unbound)."""
    try:
        import pynvml as nv
        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(device_index)
        ncpu = os.cpu_count() or 1
        words = nv.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = [64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1 and 64 * i + b < ncpu]
        if not cpus:
            return None
        allowed = os.sched_getaffinity(0)
        cpus = [c for c in cpus if c in allowed] or None
        if cpus:
            os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:
        return None
