"""ctypes view of include/seqalign.h.

Two shared objects: `core` (no CUDA; always loadable) and `cuda` (sm_100a kernels).
`cuda()` raises loudly when the library is missing — there is no silent PyTorch fallback
for the GPU path.
"""
from __future__ import annotations

import ctypes as C
from functools import lru_cache

from ..utils import build as _build

SA_OK = 0
SA_ERR_ARG = 2
SA_ERR_UNSUPPORTED = 3
SA_ERR_TOO_LARGE = 6
SA_ERR_SMALL_BUFFER = 7
SA_MODE_AUTO, SA_MODE_TILE, SA_MODE_PAIR, SA_MODE_STREAM = 0, 1, 2, 3
SA_FLAG_PAIR_MODE = 1
SA_FLAG_STREAM_MODE = 2
SA_NO_SEQ = 0xFFFFFFFF
SA_MAX_WORLD = 16
SA_NOALLOC_MAX_SEQ = 4096


class BlobHeader(C.Structure):
    _fields_ = [
        ("magic", C.c_uint32), ("version", C.c_uint32), ("total_bytes", C.c_uint32), ("flags", C.c_uint32),
        ("weights", C.c_int32 * 4),
        ("len1", C.c_uint32), ("len1_padded", C.c_uint32), ("nseq", C.c_uint32), ("nitems", C.c_uint32),
        ("off_lut", C.c_uint32), ("off_seq1", C.c_uint32), ("off_seqtab", C.c_uint32),
        ("off_items", C.c_uint32), ("off_seq2", C.c_uint32),
        ("max_rows_padded", C.c_uint32), ("max_window_bytes", C.c_uint32), ("chunk_bytes", C.c_uint32), ("nchunks", C.c_uint32), ("pad0", C.c_uint32),
        ("total_cells", C.c_uint64), ("padded_cells", C.c_uint64),
        ("world", C.c_uint32), ("rank_first_item", C.c_uint32 * (SA_MAX_WORLD + 1)),
        ("value_min", C.c_int32), ("value_range", C.c_uint32), ("off_pairs", C.c_uint32), ("npairs", C.c_uint32), ("off_tile_items", C.c_uint32),
        ("off_shared_end", C.c_uint32), ("rank_first_pair", C.c_uint32 * (SA_MAX_WORLD + 1)),
        ("rank_seq2_lo", C.c_uint32 * SA_MAX_WORLD), ("rank_seq2_hi", C.c_uint32 * SA_MAX_WORLD),
        ("reserved", C.c_uint32 * 157),
    ]


assert C.sizeof(BlobHeader) == 1024


class SeqDesc(C.Structure):
    _fields_ = [
        ("off", C.c_uint32), ("len", C.c_uint16), ("crow", C.c_uint16), ("ndiag", C.c_int32),
        ("ntiles", C.c_uint16), ("wd_log2", C.c_uint8), ("home", C.c_uint8),
    ]


assert C.sizeof(SeqDesc) == 16


class NativeError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"seqalign error {code}: {msg}")
        self.code = code


@lru_cache(maxsize=None)
def core() -> C.CDLL:
    lib = C.CDLL(str(_build.ensure_core()), mode=C.RTLD_GLOBAL)
    vp, i32p, sz = C.c_void_p, C.POINTER(C.c_int32), C.c_size_t
    lib.sa_last_error.restype = C.c_char_p
    lib.sa_parse.argtypes = [C.c_char_p, sz, C.POINTER(vp)]
    lib.sa_problem_create.argtypes = [i32p, C.c_char_p, sz, C.c_int32, C.POINTER(C.c_char_p), C.POINTER(sz), C.POINTER(vp)]
    lib.sa_problem_free.argtypes = [vp]
    lib.sa_problem_free.restype = None
    lib.sa_problem_nseq.argtypes = [vp]
    lib.sa_problem_nseq.restype = C.c_int32
    lib.sa_problem_weights.argtypes = [vp, i32p]
    lib.sa_problem_weights.restype = None
    lib.sa_problem_len1.argtypes = [vp]
    lib.sa_problem_len1.restype = sz
    lib.sa_problem_seq1.argtypes = [vp]
    lib.sa_problem_seq1.restype = C.c_void_p
    lib.sa_problem_len2.argtypes = [vp, C.c_int32]
    lib.sa_problem_len2.restype = sz
    lib.sa_problem_seq2.argtypes = [vp, C.c_int32]
    lib.sa_problem_seq2.restype = C.c_void_p
    lib.sa_problem_candidates.argtypes = [vp, C.c_int32]
    lib.sa_problem_candidates.restype = C.c_int64
    lib.sa_problem_total_candidates.argtypes = [vp]
    lib.sa_problem_total_candidates.restype = C.c_int64
    lib.sa_problem_serialize.argtypes = [vp, C.c_void_p, sz]
    lib.sa_problem_serialize.restype = sz
    lib.sa_pair_class.argtypes = [C.c_ubyte, C.c_ubyte]
    lib.sa_solve_cpu_naive.argtypes = [vp, i32p, C.c_int]
    lib.sa_solve_cpu_fast.argtypes = [vp, i32p, C.c_int]
    lib.sa_format.argtypes = [i32p, C.c_int32, C.c_void_p, sz]
    lib.sa_format.restype = sz
    lib.sa_plan_pack_into.argtypes = [vp, C.c_int, C.c_void_p, sz, C.POINTER(sz)]
    lib.sa_plan_pack_mode.argtypes = [vp, C.c_int, C.c_int, C.c_void_p, sz, C.POINTER(sz)]
    lib.sa_blob_execute_cpu.argtypes = [C.c_void_p, sz, i32p]
    lib.sa_blob_score_items_cpu.argtypes = [C.c_void_p, sz, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint64)]
    lib.sa_blob_finalize_cpu.argtypes = [C.c_void_p, sz, C.POINTER(C.c_uint64), i32p]
    lib.sa_blob_rank_ranges.argtypes = [C.POINTER(BlobHeader), C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    return lib


@lru_cache(maxsize=None)
def cuda() -> C.CDLL:
    core()  # libseqalign_cuda.so links against the core library
    path = _build.lib_path("seqalign_cuda")
    if not path.exists():
        raise RuntimeError(
            f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` or `make build`; "
            "the GPU path has no PyTorch fallback"
        )
    lib = C.CDLL(str(path))
    vp, sz = C.c_void_p, C.c_size_t
    lib.sa_cuda_scratch_bytes.argtypes = [C.c_uint32]
    lib.sa_cuda_scratch_bytes.restype = sz
    lib.sa_cuda_scratch_init.argtypes = [vp, sz, vp]
    lib.sa_cuda_solve_blob.argtypes = [vp, C.POINTER(BlobHeader), vp, vp, vp]
    lib.sa_cuda_solve_problem.argtypes = [vp, C.POINTER(C.c_int32), C.c_int]
    lib.sa_cuda_solve_blob_noalloc.argtypes = [vp, C.POINTER(BlobHeader), vp, vp]
    lib.sa_cuda_launch_count.restype = C.c_uint64
    lib.sa_session_create.argtypes = [C.c_int, C.POINTER(vp)]
    lib.sa_session_destroy.argtypes = [vp]
    lib.sa_session_destroy.restype = None
    lib.sa_session_solve.argtypes = [vp, vp, vp]  # the result pointer is passed as an address (see NativeSession.solve)
    lib.sa_session_configure.argtypes = [vp, C.c_size_t, C.c_int]
    lib.sa_session_configure.restype = None
    lib.sa_session_last_bytes.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    lib.sa_session_last_bytes.restype = None
    return lib


def check(rc: int) -> None:
    if rc != SA_OK:
        msg = core().sa_last_error()
        raise NativeError(rc, (msg or b"").decode("utf-8", "replace"))
