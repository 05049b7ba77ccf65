"""Device timing helpers (the reference has no timers at all)."""
from __future__ import annotations

import torch


class CudaTimer:
    """CUDA-event bracket on the current stream; `ms` after `stop()`+sync."""

    def __init__(self):
        self.e0 = torch.cuda.Event(enable_timing=True)
        self.e1 = torch.cuda.Event(enable_timing=True)

    def start(self):
        self.e0.record()

    def stop(self):
        self.e1.record()

    @property
    def ms(self) -> float:
        self.e1.synchronize()
        return self.e0.elapsed_time(self.e1)


class L2Flusher:
    """Writes a buffer larger than the 126 MB L2 so the next timed step starts cold."""

    def __init__(self, device, nbytes: int = 256 << 20):
        self.buf = torch.empty(nbytes, dtype=torch.uint8, device=device)
        self.v = 0

    def flush(self):
        self.v = (self.v + 1) & 0xFF
        self.buf.fill_(self.v)
