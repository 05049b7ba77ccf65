"""Planner + packer: one contiguous blob per problem (reference: main.c:110-121 split,
cudaFunctions.cu:35-61 + :189-201 uploads)."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _native as N


@dataclass
class PackedBlob:
    """Host-side packed problem. `data` is a uint8 array (numpy, or a pinned torch tensor
    when `pinned=True` was requested) holding exactly `nbytes` valid bytes."""

    data: object
    nbytes: int
    header: N.BlobHeader

    @property
    def nseq(self) -> int:
        return int(self.header.nseq)

    @property
    def nitems(self) -> int:
        return int(self.header.nitems)

    @property
    def pair_mode(self) -> bool:
        return bool(self.header.flags & N.SA_FLAG_PAIR_MODE)

    @property
    def stream_mode(self) -> bool:
        return bool(self.header.flags & N.SA_FLAG_STREAM_MODE)

    @property
    def mode(self) -> str:
        return "stream" if self.stream_mode else ("pair" if self.pair_mode else "tile")

    def numpy(self) -> np.ndarray:
        d = self.data
        return d if isinstance(d, np.ndarray) else d.numpy()

    def seq_table(self) -> np.ndarray:
        raw = self.numpy()[self.header.off_seqtab : self.header.off_seqtab + 16 * self.nseq]
        dt = np.dtype([("off", "<u4"), ("len", "<u2"), ("crow", "<u2"), ("ndiag", "<i4"),
                       ("ntiles", "<u2"), ("wd_log2", "u1"), ("home", "u1")])
        return raw.view(dt)

    def pair_records(self) -> np.ndarray:
        """STREAM mode: the self-contained 32-byte record per pair (sa_pair_rec)."""
        raw = self.numpy()[self.header.off_pairs : self.header.off_pairs + 32 * int(self.header.npairs)]
        dt = np.dtype([("a", "<u4"), ("b", "<u4"), ("off_a", "<u4"), ("off_b", "<u4"), ("len_a", "<u2"), ("len_b", "<u2"),
                       ("crow", "<u2"), ("ntiles", "<u2"), ("home", "u1"), ("pad8", "u1", 3), ("pad", "<u4")])
        return raw.view(dt)

    def items(self) -> np.ndarray:
        raw = self.numpy()[self.header.off_items : self.header.off_items + 8 * self.nitems]
        return raw.view(np.dtype([("seq", "<u4"), ("tile", "<u4")]))


MODES = {"auto": N.SA_MODE_AUTO, "tile": N.SA_MODE_TILE, "pair": N.SA_MODE_PAIR, "stream": N.SA_MODE_STREAM}


SPLIT_TILES = 0x100  # SA_MODE_SPLIT_TILES


def _mode(mode: str, split_tiles: bool) -> int:
    return MODES[mode] | (SPLIT_TILES if split_tiles else 0)


def blob_size(problem, world: int = 1, mode: str = "auto", split_tiles: bool = False) -> int:
    need = C.c_size_t()
    N.check(N.core().sa_plan_pack_mode(problem.handle, world, _mode(mode, split_tiles), None, 0, C.byref(need)))
    return int(need.value)


def plan_pack_range(problem, first: int, count: int, world: int = 1, mode: str = "auto") -> PackedBlob:
    """Plan + pack sequences [first, first + count) only (row i of that launch = sequence first + i): what the
    native session does range by range for large batches."""
    lib = N.core()
    lib.sa_plan_pack_range.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
    need = C.c_size_t()
    N.check(lib.sa_plan_pack_range(problem.handle, first, count, world, MODES[mode], None, 0, C.byref(need)))
    buf = np.empty(int(need.value), dtype=np.uint8)
    got = C.c_size_t()
    N.check(lib.sa_plan_pack_range(problem.handle, first, count, world, MODES[mode], C.c_void_p(buf.ctypes.data), int(need.value), C.byref(got)))
    hdr = N.BlobHeader.from_buffer_copy(C.string_at(buf.ctypes.data, C.sizeof(N.BlobHeader)))
    return PackedBlob(buf, int(got.value), hdr)


def plan_pack(problem, world: int = 1, pinned: bool = False, out=None, mode: str = "auto", split_tiles: bool = False) -> PackedBlob:
    """Plan the tiling, partition over `world` GPUs and pack. Raises NativeError with
    code SA_ERR_UNSUPPORTED when the input is outside the GPU executor's limits.
    mode: 'auto' | 'tile' (any input) | 'pair' (u16x2 throughput mode, single GPU) | 'stream';
    split_tiles: deal TILE items tile by tile so that single sequences are split over GPUs."""
    got = C.c_size_t()
    if out is not None:
        # one pass straight into the caller's (pinned) buffer; only a buffer that turns out too small costs a sizing pass
        buf, ptr = out, out.data_ptr()
        rc = N.core().sa_plan_pack_mode(problem.handle, world, _mode(mode, split_tiles), C.c_void_p(ptr), int(buf.numel()), C.byref(got))
        if rc == N.SA_ERR_SMALL_BUFFER:
            raise ValueError("provided buffer too small")
        N.check(rc)
    else:
        need = blob_size(problem, world, mode, split_tiles)
        if pinned:
            import torch

            buf = torch.empty(need, dtype=torch.uint8, pin_memory=True)
            ptr = buf.data_ptr()
        else:
            buf = np.empty(need, dtype=np.uint8)
            ptr = buf.ctypes.data
        N.check(N.core().sa_plan_pack_mode(problem.handle, world, _mode(mode, split_tiles), C.c_void_p(ptr), need, C.byref(got)))
    hdr = N.BlobHeader.from_buffer_copy(C.string_at(ptr, C.sizeof(N.BlobHeader)))
    return PackedBlob(buf, int(got.value), hdr)


def execute_blob_cpu(blob: PackedBlob) -> np.ndarray:
    """Walk the blob on the CPU with the GPU kernel's tile logic -> int32[nseq, 3]."""
    arr = blob.numpy()
    out = np.zeros((max(blob.nseq, 1), 4), dtype=np.int32)
    N.check(N.core().sa_blob_execute_cpu(C.c_void_p(arr.ctypes.data), blob.nbytes, out.ctypes.data_as(C.POINTER(C.c_int32))))
    return out[: blob.nseq, :3].copy()
