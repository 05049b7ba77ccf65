#!/usr/bin/env python
"""One-GPU probe of the scatter path: plays rank 1 of a world of two with BOTH symmetric buffers on
the same GPU (the header packet is written by the host, 'rank 0' never runs), so that the cost of
scoring from freshly staged chunks can be separated from anything NVLink does.

    python benchmarks/staging_probe.py [--scale 512]

Prints, for the fabric kernel as rank 1: fresh (pull + score) vs resident (score only) times and the
in-kernel timeline of each."""
from __future__ import annotations

import argparse
import ctypes as C
import json
import statistics
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import numpy as np  # noqa: E402
import torch  # noqa: E402

from mpi_openmp_cuda_b200.models.synthetic import scaled_input3  # noqa: E402
from mpi_openmp_cuda_b200.ops import _native as N  # noqa: E402
from mpi_openmp_cuda_b200.ops import plan  # noqa: E402
from mpi_openmp_cuda_b200.parallel.fabric import FabricDesc, FabricLayout  # noqa: E402
from mpi_openmp_cuda_b200.utils.timing import L2Flusher  # noqa: E402

NAMES = {1: "header_known", 17: "first_chunk_staged", 5: "all_chunks_staged", 2: "items_done", 16: "last_cta_exit",
         8: "q1", 9: "q2", 10: "q3", 18: "first_warp_idle", 19: "last_warp_idle"}


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--scale", type=int, default=512, help="input3 x SCALE per GPU, world of 2")
    ap.add_argument("--mode", default="auto")
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--no-flush", action="store_true")
    a = ap.parse_args()
    lib = N.cuda()
    lib.sa_fabric_layout.argtypes = [C.c_uint32, C.c_size_t, C.POINTER(FabricLayout)]
    lib.sa_cuda_fabric_solve.argtypes = [C.POINTER(FabricDesc), C.POINTER(N.BlobHeader), C.c_void_p]
    p = scaled_input3(a.scale * 2, ROOT / "data" / "input3.txt", seed=1)
    blob = plan.plan_pack(p, 2, mode=a.mode)
    lay = FabricLayout()
    N.check(lib.sa_fabric_layout(max(blob.nseq, 64), blob.nbytes + (1 << 20), C.byref(lay)))
    bufs = [torch.zeros(int(lay.total_bytes), dtype=torch.uint8, device="cuda:0") for _ in range(2)]
    bufs[0][lay.off_blob : lay.off_blob + blob.nbytes].copy_(torch.from_numpy(blob.numpy()[: blob.nbytes]))
    hdr_words = np.frombuffer(bytes(blob.header), dtype=np.uint32).astype(np.uint64)
    fl = L2Flusher("cuda:0")
    d = FabricDesc()
    d.rank, d.world, d.use_multicast, d.timeout_ms = 1, 2, 0, 2000
    d.peer_base[0], d.peer_base[1] = bufs[0].data_ptr(), bufs[1].data_ptr()
    d.layout = lay
    tl = bufs[1][lay.off_ctrl + 64 : lay.off_ctrl + 64 + 8 * 24].view(torch.int64)
    epoch = 0

    def run(resident: bool):
        nonlocal epoch
        if not resident:
            epoch += 1
            pkt = torch.from_numpy(((np.uint64(epoch) << np.uint64(32)) | hdr_words).view(np.int64).copy()).cuda()
            bufs[1][lay.off_hdr_ll : lay.off_hdr_ll + 8 * len(hdr_words)].view(torch.int64).copy_(pkt)
        d.epoch = epoch
        d.pad0 = 1 | (2 if resident else 0)
        if not a.no_flush:
            fl.flush()
        tl.zero_()
        tl[0] = torch.iinfo(torch.int64).max
        tl[17] = torch.iinfo(torch.int64).max
        tl[18] = torch.iinfo(torch.int64).max
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        N.check(lib.sa_cuda_fabric_solve(C.byref(d), None, None))
        e1.record()
        e1.synchronize()
        t = tl.cpu().tolist()
        big = torch.iinfo(torch.int64).max
        err = int(bufs[1][lay.off_ctrl + 12 : lay.off_ctrl + 16].view(torch.int32).item())
        tlr = {n: (None if t[i] in (0, big) else round((t[i] - t[0]) / 1e3, 1)) for i, n in NAMES.items()}
        tlr["items_that_waited"], tlr["wait_us_total"] = t[6], round(t[7] / 1e3, 1)
        tlr["longest_item_who"] = {"start_us": t[23] >> 40, "item": (t[23] >> 16) & 0xFFFFFF, "warp": (t[23] >> 8) & 0xFF, "odd_cta": (t[23] >> 1) & 1, "copied_before": t[23] & 1}
        tlr["longest_item_us"], tlr["longest_item_last_quarter_us"], tlr["items_executed"], tlr["rows_walked"] = round(t[20] / 1e3, 1), round(t[21] / 1e3, 1), t[3], t[22]
        return e0.elapsed_time(e1) * 1e3, tlr, err

    out = {"config": f"input3 x{a.scale} per GPU, world 2, rank 1 alone on one GPU", "mode": blob.mode, "nseq": blob.nseq,
           "blob_bytes": blob.nbytes, "chunk_bytes": int(blob.header.chunk_bytes)}
    for name, resident in (("fresh", False), ("resident", True), ("fresh_again", False)):
        ts, last, err = [], None, 0
        for _ in range(a.iters):
            us, last, err = run(resident)
            ts.append(us)
        out[name] = {"event_us": round(statistics.median(ts), 1), "min_us": round(min(ts), 1), "timeline": last, "error_word": err}
    print(json.dumps(out))
    return 0


if __name__ == "__main__":
    sys.exit(main())
