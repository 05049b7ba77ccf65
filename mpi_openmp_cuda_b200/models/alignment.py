"""The flagship "model": an exhaustive single-gap alignment solver for one GPU.

`AlignmentSolver.solve(problem)` is the call a user makes; it is the whole reference
pipeline after parsing (main.c:110-197 + cudaFunctions.cu:178-242): plan + pack on the host
into pinned memory, one H2D copy, one kernel launch, one D2H copy of the result table.
"""
from __future__ import annotations

import numpy as np
import torch

from ..ops import _native as N
from ..ops import oracle, plan
from ..ops.score import GpuExecutor, NativeSession


class AlignmentSolver:
    def __init__(self, device: str | torch.device | int = "cuda:0"):
        self.exec = GpuExecutor(device)
        self.device = self.exec.device
        self.session = NativeSession(self.device.index or 0)
        self._pin_in = None
        self._pin_out = None

    # -- pinned staging
    def _pinned_in(self, nbytes: int) -> torch.Tensor:
        if self._pin_in is None or self._pin_in.numel() < nbytes:
            self._pin_in = torch.empty(max(nbytes, 1 << 16), dtype=torch.uint8, pin_memory=True)
        return self._pin_in

    def _pinned_out(self, nseq: int) -> torch.Tensor:
        if self._pin_out is None or self._pin_out.shape[0] < nseq:
            self._pin_out = torch.empty((max(nseq, 64), 4), dtype=torch.int32, pin_memory=True)
        return self._pin_out

    def pack(self, problem, mode: str = "auto") -> plan.PackedBlob:
        """Plan + pack straight into this solver's pinned staging buffer."""
        try:  # one pass into the staging buffer we already have; size + grow only when it is too small
            return plan.plan_pack(problem, 1, out=self._pinned_in(1), mode=mode)
        except ValueError:
            need = plan.blob_size(problem, 1, mode)
            return plan.plan_pack(problem, 1, out=self._pinned_in(need), mode=mode)

    def solve_packed(self, blob: plan.PackedBlob, sync: bool = True) -> np.ndarray:
        """H2D -> kernel -> D2H for an already packed problem. Returns int32[nseq,3]."""
        nseq = blob.nseq
        if nseq == 0:
            return np.zeros((0, 3), dtype=np.int32)
        with torch.cuda.device(self.device):
            dev_out = self.exec.solve_blob(blob)
            host = self._pinned_out(nseq)
            host[:nseq].copy_(dev_out, non_blocking=True)
            if sync:
                torch.cuda.current_stream().synchronize()
        return host[:nseq, :3].numpy()

    def solve(self, problem) -> np.ndarray:
        """The user-facing call: one native session solve (pack, H2D, kernel, D2H, sync)."""
        try:
            return self.session.solve(problem)
        except N.NativeError as e:
            if e.code == N.SA_ERR_UNSUPPORTED:  # outside the GPU executor's limits
                return oracle.solve_cpu(problem, "fast")
            raise

    def solve_via_torch(self, problem) -> np.ndarray:
        """Same result through torch-owned buffers on the current torch stream (pack() + solve_packed())."""
        return self.solve_packed(self.pack(problem)).copy()


_default: dict[str, AlignmentSolver] = {}


def solve(problem, device: str | None = None) -> np.ndarray:
    """GPU when available (no silent fallback once a GPU is present), CPU oracle otherwise."""
    if device is None:
        device = "cuda:0" if torch.cuda.is_available() else "cpu"
    if str(device).startswith("cpu"):
        return oracle.solve_cpu(problem, "fast")
    key = str(device)
    if key not in _default:
        _default[key] = AlignmentSolver(device)
    return _default[key].solve(problem)
