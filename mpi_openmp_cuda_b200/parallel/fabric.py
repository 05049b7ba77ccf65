"""Bootstrap of the GPU fabric: one process per GPU, a symmetric buffer per rank mapped into
every peer (NVLink 5 / NVSwitch), plus the NVLS multicast alias when the platform has one.

torch.distributed (+ torch symmetric memory) is used ONLY here, to exchange the mappings once.
The data path runs inside `sa_score_fabric_kernel` (csrc/cuda/align_kernel.cu) and replaces the
reference's MPI_Bcast/Scatter/Gather (main.c:149-152,174,195-197): rank 0 multicasts the 1 KB
header as a flag-in-data packet, every other rank pulls the shared table/Seq1 and its OWN rows
from rank 0's buffer (a scatter: nothing is sent that is not consumed), finished rows are stored
straight into rank 0's table as self-validating words, split sequences reduce with atomicMax.sys.
"""
from __future__ import annotations

import ctypes as C
import os

import torch
import torch.distributed as dist

from ..ops import _native as N


class FabricLayout(C.Structure):
    _fields_ = [("off_ctrl", C.c_uint32), ("off_hdr_ll", C.c_uint32), ("off_key", C.c_uint32), ("off_cnt", C.c_uint32),
                ("off_out", C.c_uint32), ("off_out_ll", C.c_uint32), ("off_blob", C.c_uint32), ("max_seq", C.c_uint32),
                ("max_blob_bytes", C.c_uint64), ("total_bytes", C.c_uint64)]


class FabricDesc(C.Structure):
    _fields_ = [("rank", C.c_uint32), ("world", C.c_uint32), ("epoch", C.c_uint32), ("use_multicast", C.c_uint32),
                ("timeout_ms", C.c_uint32), ("pad0", C.c_uint32),
                ("peer_base", C.c_void_p * N.SA_MAX_WORLD), ("mc_base", C.c_void_p), ("layout", FabricLayout)]


def init_distributed(backend: str | None = None) -> tuple[int, int, int]:
    """Join the job described by RANK / WORLD_SIZE / LOCAL_RANK / MASTER_* (torchrun or our launcher)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if backend == "nccl":
        torch.cuda.set_device(local)
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29531")
        kw = {"device_id": torch.device("cuda", local)} if backend == "nccl" else {}
        rdzv = os.environ.get("SA_RDZV_FILE")  # our launcher: a file-store rendezvous, no TCP port to collide on
        if rdzv:
            kw["init_method"] = f"file://{rdzv}"
        dist.init_process_group(backend, rank=rank, world_size=world, **kw)
    return rank, world, local


class Fabric:
    """Symmetric buffer + peer mappings for `world` GPUs of one node."""

    def __init__(self, max_seq: int = 1 << 16, max_blob_bytes: int = 96 << 20, use_multicast: bool | None = None,
                 timeout_ms: int = 10000, group=None):
        import torch.distributed._symmetric_memory as symm

        if not dist.is_initialized():
            raise RuntimeError("Fabric needs torch.distributed (see init_distributed)")
        self.group = group or dist.group.WORLD
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        if self.world > N.SA_MAX_WORLD:
            raise ValueError(f"at most {N.SA_MAX_WORLD} GPUs")
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.lib = N.cuda()
        self.lib.sa_fabric_layout.argtypes = [C.c_uint32, C.c_size_t, C.POINTER(FabricLayout)]
        self.lib.sa_cuda_fabric_solve.argtypes = [C.POINTER(FabricDesc), C.POINTER(N.BlobHeader), C.c_void_p]
        self.layout = FabricLayout()
        N.check(self.lib.sa_fabric_layout(max_seq, max_blob_bytes, C.byref(self.layout)))
        self.buf = symm.empty(int(self.layout.total_bytes), dtype=torch.uint8, device=self.device)
        self.buf.zero_()
        torch.cuda.synchronize()
        self.hdl = symm.rendezvous(self.buf, self.group)
        self.peer_ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        mc = int(self.hdl.multicast_ptr) if self.hdl.has_multicast_support else 0
        env = os.environ.get("SA_FABRIC_MULTICAST")
        if use_multicast is None:
            use_multicast = (env != "0") if env is not None else True
        self.mc_ptr = mc
        self.use_multicast = bool(use_multicast and mc != 0 and self.world > 1)
        self.timeout_ms = timeout_ms
        self.timeline = os.environ.get("SA_FABRIC_TIMELINE") == "1"
        self.epoch = 0
        d = FabricDesc()
        d.rank, d.world = self.rank, self.world
        d.use_multicast = 1 if self.use_multicast else 0
        d.timeout_ms = self.timeout_ms
        d.pad0 = 1 if self.timeline else 0
        for r in range(self.world):
            d.peer_base[r] = self.peer_ptrs[r]
        d.mc_base = self.mc_ptr or None
        d.layout = self.layout
        self._desc = d
        self.hdl.barrier()  # every rank's buffer is zeroed and mapped before anyone launches

    # views into this rank's own buffer
    def region(self, off: int, nbytes: int) -> torch.Tensor:
        return self.buf[off : off + nbytes]

    @property
    def blob_region(self) -> torch.Tensor:
        return self.buf[self.layout.off_blob : self.layout.off_blob + self.layout.max_blob_bytes]

    def out_region(self, nseq: int) -> torch.Tensor:
        return self.buf[self.layout.off_out : self.layout.off_out + 16 * nseq].view(torch.int32).view(nseq, 4)

    def error_word(self) -> int:
        return int(self.buf[self.layout.off_ctrl + 12 : self.layout.off_ctrl + 16].view(torch.int32).item())

    def read_timeline(self) -> dict:
        """After a synchronised launch with SA_FABRIC_TIMELINE=1: microseconds since this rank's first CTA started."""
        t = self.buf[self.layout.off_ctrl + 64 : self.layout.off_ctrl + 64 + 8 * 18].view(torch.int64).cpu().tolist()
        names = {1: "header_known", 17: "first_chunk_staged", 5: "all_chunks_staged", 11: "first_item_known",
                 12: "first_tma_landed", 13: "first_scan_done", 15: "first_result_posted", 2: "items_done",
                 4: "all_results_validated", 16: "last_cta_exit"}
        big = torch.iinfo(torch.int64).max
        out = {"start": 0.0}
        out.update({n: (None if t[i] in (0, big) else round((t[i] - t[0]) / 1e3, 2)) for i, n in names.items()})
        out["items_that_waited"] = t[6]
        out["wait_us_total"] = round(t[7] / 1e3, 1)
        out["quartiles_handed_out"] = [None if v == 0 else round((v - t[0]) / 1e3, 1) for v in t[8:11]]
        return out

    # -- the two primitives in isolation (benchmarks/fabric_sweep.py)
    def bcast(self, nbytes: int, ctas: int = 0) -> None:
        """Replicate rank 0's blob region [0, nbytes) into every peer's blob region (collective)."""
        self.lib.sa_cuda_fabric_bcast.argtypes = [C.POINTER(FabricDesc), C.c_uint64, C.c_int, C.c_void_p]
        self.epoch += 1
        self._desc.epoch = self.epoch
        stream = torch.cuda.current_stream(self.device).cuda_stream
        N.check(self.lib.sa_cuda_fabric_bcast(C.byref(self._desc), int(nbytes), int(ctas), C.c_void_p(stream)))

    def pull(self, nbytes: int, ctas: int = 0) -> None:
        """Every rank but 0 copies rank 0's blob region [0, nbytes) into its own (collective; rank 0's bytes must be in place)."""
        self.lib.sa_cuda_fabric_pull.argtypes = [C.POINTER(FabricDesc), C.c_uint64, C.c_int, C.c_void_p]
        stream = torch.cuda.current_stream(self.device).cuda_stream
        N.check(self.lib.sa_cuda_fabric_pull(C.byref(self._desc), int(nbytes), int(ctas), C.c_void_p(stream)))

    def argmax(self, keys: torch.Tensor, ctas: int = 0) -> None:
        """Max-fold this rank's packed u64 keys (int64 tensor, non-negative) into rank 0's blob region."""
        self.lib.sa_cuda_fabric_argmax.argtypes = [C.POINTER(FabricDesc), C.c_void_p, C.c_uint64, C.c_int, C.c_void_p]
        self.epoch += 1
        self._desc.epoch = self.epoch
        stream = torch.cuda.current_stream(self.device).cuda_stream
        N.check(self.lib.sa_cuda_fabric_argmax(C.byref(self._desc), C.c_void_p(keys.data_ptr()), int(keys.numel()),
                                               int(ctas), C.c_void_p(stream)))

    def launch(self, header: N.BlobHeader | None, resident: bool = False) -> None:
        """Enqueue this rank's kernel of the next solve on the current stream.
        resident=True re-solves the blob every rank already holds from the previous launch
        (no broadcast; every rank must pass the same value)."""
        if not resident:
            self.epoch += 1
        d = self._desc
        d.epoch = self.epoch
        d.pad0 = (1 if self.timeline else 0) | (2 if resident else 0)
        if self.timeline:  # slot 0 is a running minimum: re-arm it (and the maxima) before every launch
            t = self.buf[self.layout.off_ctrl + 64 : self.layout.off_ctrl + 64 + 8 * 18].view(torch.int64)
            t.zero_()
            t[0] = torch.iinfo(torch.int64).max
            t[17] = torch.iinfo(torch.int64).max
        stream = torch.cuda.current_stream(self.device).cuda_stream
        hdr = C.byref(header) if header is not None else None
        with torch.cuda.device(self.device):
            N.check(self.lib.sa_cuda_fabric_solve(C.byref(d), hdr, C.c_void_p(stream)))
