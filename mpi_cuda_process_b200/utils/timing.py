"""Device timing helpers.  The reference contains no timing code of any kind."""
from __future__ import annotations

import time
from typing import Callable, List, Optional

import torch


class L2Flusher:
    """Writes a buffer larger than the 126 MB L2 between timed iterations."""

    def __init__(self, device, nbytes: int = 256 << 20):
        self.buf = torch.empty(nbytes // 4, dtype=torch.float32, device=device)
        self._v = 0.0

    def flush(self) -> None:
        self._v += 1.0
        self.buf.fill_(self._v)


def time_cuda(fn: Callable[[], None], iters: int = 10, warmup: int = 3, flush: Optional[L2Flusher] = None) -> List[float]:
    """Per-iteration device times in ms (CUDA events on the current stream, sync on both sides)."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    out = []
    for _ in range(iters):
        if flush is not None:
            flush.flush()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        out.append(e0.elapsed_time(e1))
    return out


class WallTimer:
    def __enter__(self):
        self.t0 = time.perf_counter()
        return self

    def __exit__(self, *a):
        self.ms = (time.perf_counter() - self.t0) * 1e3
        return False
