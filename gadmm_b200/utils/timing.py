"""Device-side timing helpers (CUDA events on the launching stream, after warm-up, synchronised on
both sides -- never numbers taken under a profiler)."""
from __future__ import annotations

import torch


def cuda_time_ms(fn, warmup: int = 3, reps: int = 5, sync_device=None) -> float:
    """Median device time of ``fn()`` in milliseconds."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize(sync_device)
    out = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize(sync_device)
        out.append(e0.elapsed_time(e1))
    out.sort()
    return out[len(out) // 2]


def flush_l2(device=None, nbytes: int = 256 << 20):
    """Write a buffer larger than the 126 MB L2 (for kernels whose inputs would otherwise stay
    resident between timed repetitions)."""
    buf = torch.empty(nbytes, dtype=torch.uint8, device=device or torch.device("cuda"))
    buf.fill_(1)
    return buf
