"""Multi-GPU solver: the reference's master/worker run (main.c:46-244) as one kernel per GPU.

Rank 0 is the only rank that ever sees the problem (as in the reference, where only the root
reads stdin): it plans for `world` GPUs, packs, copies the blob into its own symmetric buffer
and launches. Every other rank just launches — its kernel receives the data over NVLink from
rank 0's kernel, scores its share and writes results into rank 0's table.
"""
from __future__ import annotations

import numpy as np
import torch

from ..ops import plan
from .fabric import Fabric


class DistributedSolver:
    def __init__(self, fabric: Fabric):
        self.fabric = fabric
        self._pin_in = None
        self._pin_out = None
        self._pin_err = torch.zeros(1, dtype=torch.int32, pin_memory=True) if torch.cuda.is_available() else None
        self._err_fresh = False

    @property
    def rank(self) -> int:
        return self.fabric.rank

    def _pinned_in(self, nbytes: int) -> torch.Tensor:
        if self._pin_in is None or self._pin_in.numel() < nbytes:
            self._pin_in = torch.empty(max(nbytes, 1 << 16), dtype=torch.uint8, pin_memory=True)
        return self._pin_in

    def pack(self, problem) -> plan.PackedBlob:
        try:  # one pass into the staging buffer we already have; size + grow only when it is too small
            return plan.plan_pack(problem, self.fabric.world, out=self._pinned_in(1))
        except ValueError:
            need = plan.blob_size(problem, self.fabric.world)
            return plan.plan_pack(problem, self.fabric.world, out=self._pinned_in(need))

    def launch_packed(self, blob: plan.PackedBlob | None) -> None:
        """All ranks call this once per solve; only rank 0 passes a blob."""
        f = self.fabric
        if f.rank == 0:
            if blob is None:
                raise ValueError("rank 0 must provide the packed problem")
            if blob.nbytes > f.layout.max_blob_bytes or blob.nseq > f.layout.max_seq:
                raise ValueError("problem exceeds the fabric's capacity; create the Fabric with larger limits")
            src = blob.data if isinstance(blob.data, torch.Tensor) else torch.from_numpy(blob.data)
            f.blob_region[: blob.nbytes].copy_(src[: blob.nbytes], non_blocking=True)
            f.launch(blob.header)
        else:
            f.launch(None)

    def fetch(self, nseq: int, sync: bool = True, with_error: bool = False) -> np.ndarray | None:
        """Rank 0: D2H of the result table (after its kernel, which ends only when all ranks delivered)."""
        f = self.fabric
        # this rank's error word rides along with the results: one synchronisation per solve, not two
        if with_error and self._pin_err is not None:
            off = f.layout.off_ctrl + 12
            self._pin_err.copy_(f.buf[off : off + 4].view(torch.int32), non_blocking=True)
            self._err_fresh = sync
        if f.rank != 0:
            if sync:
                torch.cuda.current_stream().synchronize()
            return None
        if self._pin_out is None or self._pin_out.shape[0] < nseq:
            self._pin_out = torch.empty((max(nseq, 64), 4), dtype=torch.int32, pin_memory=True)
        self._pin_out[:nseq].copy_(f.out_region(nseq), non_blocking=True)
        if sync:
            torch.cuda.current_stream().synchronize()
        return self._pin_out[:nseq, :3].numpy()

    def check(self) -> None:
        if self._err_fresh:  # fetched together with the results of the solve that just ended
            self._err_fresh = False
            err = int(self._pin_err[0])
        else:
            err = self.fabric.error_word()
        if err:
            raise RuntimeError(f"fabric wait timed out on rank {self.rank} (code {err}): a peer GPU did not deliver")

    def solve(self, problem=None) -> np.ndarray | None:
        """Collective: every rank calls solve(); rank 0 passes the problem and gets int32[nseq,3].
        An input outside the GPU executor's limits (Seq2 > 8192 letters, more than five distinct non-letter
        bytes, very large weights) is solved by the CPU oracle on rank 0 — as `final --gpus 1` does — while
        the fabric runs an empty problem, so that the other ranks' kernels return at once instead of
        waiting for a header that would never come."""
        if self.rank == 0:
            from ..ops import _native as N
            from ..ops import oracle

            fallback = None
            try:
                blob = self.pack(problem)
                if blob.nbytes > self.fabric.layout.max_blob_bytes or blob.nseq > self.fabric.layout.max_seq:
                    raise N.NativeError(N.SA_ERR_UNSUPPORTED, "problem exceeds the fabric's capacity")
            except N.NativeError as e:
                if e.code != N.SA_ERR_UNSUPPORTED:
                    raise
                fallback = oracle.solve_cpu(problem, "fast")
                from ..utils.parser import Problem

                blob = self.pack(Problem.from_parts((1, 1, 1, 1), b"A", []))
            self.launch_packed(blob)
            res = self.fetch(blob.nseq, with_error=True)
            self.check()
            return fallback if fallback is not None else res.copy()
        self.launch_packed(None)
        self.fetch(0, with_error=True)
        self.check()
        return None
