"""GPU executor binding: one kernel launch per problem on the current CUDA stream.

Replaces the reference's `send_divided_Seq2_To_Cuda` host loop (cudaFunctions.cu:178-242:
per-row cudaMalloc + cudaGetDeviceProperties + launch + cudaDeviceSynchronize) with
persistent buffers and a single persistent kernel (csrc/cuda/align_kernel.cu).
There is no PyTorch fallback: if libseqalign_cuda.so is missing this raises.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _native as N
from .plan import PackedBlob


def launch_count() -> int:
    """Kernels launched by libseqalign_cuda.so in this process (bench `gpu_launches`)."""
    return int(N.cuda().sa_cuda_launch_count())


class GpuExecutor:
    """Owns the device-side blob / scratch / result buffers of one GPU and reuses them
    across calls (grown geometrically, never shrunk)."""

    def __init__(self, device: torch.device | str | int = "cuda:0"):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise ValueError("GpuExecutor needs a CUDA device")
        self.lib = N.cuda()
        self._blob = None
        self._out = None
        self._scratch = None
        self._scratch_nseq = 0

    def _ensure(self, nbytes: int, nseq: int) -> None:
        if self._blob is None or self._blob.numel() < nbytes:
            cap = max(nbytes, 1 << 16)
            if self._blob is not None:
                cap = max(cap, 2 * self._blob.numel())
            self._blob = torch.empty(cap, dtype=torch.uint8, device=self.device)
        if self._out is None or self._out.shape[0] < nseq:
            cap = max(nseq, 64 if self._out is None else 2 * self._out.shape[0])
            self._out = torch.empty((cap, 4), dtype=torch.int32, device=self.device)
        if self._scratch is None or self._scratch_nseq < nseq:
            cap = max(nseq, 64 if self._scratch is None else 2 * self._scratch_nseq)
            # zero-filled once; the kernel hands it back zeroed after every launch
            self._scratch = torch.zeros(int(self.lib.sa_cuda_scratch_bytes(cap)), dtype=torch.uint8, device=self.device)
            self._scratch_nseq = cap

    def upload(self, blob: PackedBlob) -> None:
        """Async H2D of the packed problem on the current stream (pinned source => truly async)."""
        self._ensure(blob.nbytes, blob.nseq)
        src = blob.data
        if not isinstance(src, torch.Tensor):
            src = torch.from_numpy(src)
        self._blob[: blob.nbytes].copy_(src[: blob.nbytes], non_blocking=True)

    def run(self, header: N.BlobHeader) -> torch.Tensor:
        """Launch on the current stream over the blob already in `self._blob`.
        Returns the device tensor int32[nseq, 4] = (score, n, k, 0)."""
        nseq = int(header.nseq)
        self._ensure(int(header.total_bytes), nseq)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        with torch.cuda.device(self.device):
            # scratch sized for capacity >= nseq: layout depends on nseq, so carve with this nseq
            N.check(self.lib.sa_cuda_solve_blob(
                C.c_void_p(self._blob.data_ptr()), C.byref(header), C.c_void_p(self._out.data_ptr()),
                C.c_void_p(self._scratch.data_ptr()), C.c_void_p(stream)))
        return self._out[:nseq]

    def solve_blob(self, blob: PackedBlob) -> torch.Tensor:
        self.upload(blob)
        return self.run(blob.header)


class NativeSession:
    """`sa_session`: the whole per-solve host path (plan + pack into pinned memory, H2D, launch,
    D2H, synchronise) in ONE native call on a private stream — no Python between the steps."""

    def __init__(self, device_index: int = 0):
        self.lib = N.cuda()
        self._h = C.c_void_p()
        self._stage = {}
        N.check(self.lib.sa_session_create(int(device_index), C.byref(self._h)))

    def solve(self, problem):
        import numpy as np

        n = problem.nseq
        if n <= 4096:
            # small problems: a staging array per size whose address is looked up once — asking numpy for an
            # array's address (ndarray.ctypes / __array_interface__) costs 3-5 us per call, a tenth of a small solve
            hit = self._stage.get(n)
            if hit is None:
                buf = np.empty((max(n, 1), 3), dtype=np.int32)
                hit = self._stage[n] = (buf, buf.ctypes.data)
                if len(self._stage) > 64:
                    self._stage.pop(next(iter(self._stage)))
            N.check(self.lib.sa_session_solve(self._h, problem.handle, hit[1]))
            return hit[0][:n].copy()
        out = np.empty((n, 3), dtype=np.int32)
        N.check(self.lib.sa_session_solve(self._h, problem.handle, out.ctypes.data))
        return out

    def configure(self, zero_copy_max_bytes: int = 0, direct_out: bool = True) -> None:
        self.lib.sa_session_configure(self._h, int(zero_copy_max_bytes), 1 if direct_out else 0)

    def set_pipeline(self, min_seq: int) -> None:
        """Batches of at least `min_seq` sequences are packed / copied / launched range by range (0 = never)."""
        self.lib.sa_session_set_pipeline.argtypes = [C.c_void_p, C.c_size_t]
        self.lib.sa_session_set_pipeline(self._h, int(min_seq))

    def last_launches(self) -> int:
        self.lib.sa_session_last_launches.restype = C.c_uint32
        self.lib.sa_session_last_launches.argtypes = [C.c_void_p]
        return int(self.lib.sa_session_last_launches(self._h))

    def last_bytes(self) -> tuple[int, int]:
        a, b = C.c_uint64(), C.c_uint64()
        self.lib.sa_session_last_bytes(self._h, C.byref(a), C.byref(b))
        return int(a.value), int(b.value)

    def close(self):
        if self._h is not None and self._h.value:
            self.lib.sa_session_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
