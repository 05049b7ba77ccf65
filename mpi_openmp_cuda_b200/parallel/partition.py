"""Work partition over GPUs (reference: main.c:110-121 `rows = N / P` + remainder on root).

The partition itself is computed by the native planner (`sa_plan_pack(problem, world)`) and
travels inside the blob; this module only exposes it for inspection and tests."""
from __future__ import annotations

import numpy as np

from ..ops import plan


def rank_item_ranges(blob: plan.PackedBlob) -> list[tuple[int, int]]:
    h = blob.header
    return [(int(h.rank_first_item[r]), int(h.rank_first_item[r + 1])) for r in range(int(h.world))]


def rank_loads(blob: plan.PackedBlob) -> np.ndarray:
    """Padded cells assigned to every rank."""
    tab, items = blob.seq_table(), blob.items()
    wd = tab["wd_log2"].astype(np.int64)
    cost = (256 << wd) * (16 >> wd) * tab["crow"].astype(np.int64)
    return np.array([int(cost[items[a:b]["seq"]].sum()) for a, b in rank_item_ranges(blob)], dtype=np.int64)


def reference_partition(nseq: int, nprocs: int) -> tuple[int, int]:
    """The reference's static split (main.c:110-119): (rows per rank, remainder on root)."""
    if nprocs < nseq:
        rows = nseq // nprocs
        return rows, nseq - nprocs * rows
    return 1, 0
