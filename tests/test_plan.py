"""Planner/packer + CPU walk of the packed blob (the GPU kernel's tile logic on the host)."""
import numpy as np
import pytest

from mpi_openmp_cuda_b200 import Problem, format_results
from mpi_openmp_cuda_b200.models.synthetic import random_problem
from mpi_openmp_cuda_b200.ops import plan
from mpi_openmp_cuda_b200.ops._native import NativeError, SA_ERR_UNSUPPORTED
from mpi_openmp_cuda_b200.ops.oracle import solve_cpu


@pytest.mark.parametrize("i", [1, 2, 3, 4, 5, 6])
@pytest.mark.parametrize("world", [1, 2, 8])
def test_blob_walk_matches_golden(shipped_inputs, i, world):
    path, gold = shipped_inputs[i]
    blob = plan.plan_pack(Problem.from_file(path), world)
    assert format_results(plan.execute_blob_cpu(blob)) == gold.read_bytes()


def test_header_and_alignment(shipped_inputs):
    blob = plan.plan_pack(Problem.from_file(shipped_inputs[3][0]), 4)
    h = blob.header
    assert h.total_bytes == blob.nbytes and h.nseq == 32 and h.world == 4
    assert h.off_lut % 16 == 0 and h.off_seq1 % 16 == 0 and h.off_seq2 % 16 == 0
    tab = blob.seq_table()
    assert all(int(o) % 16 == 0 for o in tab["off"])
    assert int(tab["ntiles"].sum()) == h.nitems
    first = list(h.rank_first_item)[:5]
    assert first[0] == 0 and first[4] == h.nitems and first == sorted(first)


def test_every_item_appears_once(shipped_inputs):
    blob = plan.plan_pack(Problem.from_file(shipped_inputs[4][0]), 8)
    items = blob.items()
    seen = set((int(s), int(t)) for s, t in zip(items["seq"], items["tile"]))
    assert len(seen) == blob.nitems
    tab = blob.seq_table()
    for s in range(blob.nseq):
        for t in range(int(tab["ntiles"][s])):
            assert (s, t) in seen


def test_partition_balance_large():
    p = random_problem(512, len1=900, len2_range=(50, 700), seed=3)
    blob = plan.plan_pack(p, 8)
    h, tab, items = blob.header, blob.seq_table(), blob.items()
    cost = (256 << tab["wd_log2"].astype(np.int64)) * (16 >> tab["wd_log2"].astype(np.int64)) * tab["crow"].astype(np.int64)
    loads = []
    for r in range(8):
        sl = items[h.rank_first_item[r] : h.rank_first_item[r + 1]]
        loads.append(int(cost[sl["seq"]].sum()))
        # whole sequences stay on one rank when sequences are plentiful
        assert all(int(tab["home"][s]) == r for s in sl["seq"])
    assert max(loads) <= 1.1 * (sum(loads) / 8)


def test_scarce_sequences_are_split_by_tile(shipped_inputs):
    blob = plan.plan_pack(Problem.from_file(shipped_inputs[2][0]), 8)  # one sequence
    assert blob.nitems >= 1
    p = Problem.from_parts((1, 1, 1, 1), "A" * 2900, ["C" * 100])
    blob = plan.plan_pack(p, 4)
    h = blob.header
    counts = [h.rank_first_item[r + 1] - h.rank_first_item[r] for r in range(4)]
    assert sum(counts) == blob.nitems and max(counts) - min(counts) <= 1


EDGE = [
    ("ABCDEFGH", ["A"]),             # len2 = 1: only k = 0 exists
    ("ABCDEFGH", ["AB", "ABC"]),     # len2 in {2,3}: the reference drops classes here (Q8)
    ("ABCDEFGH", ["ABCDEFGH"]),      # equal length
    ("ABCDEFGH", ["ABCDEFG"]),       # len1 - 1: a single offset
    ("ABC", ["ABCD", "AB"]),         # longer than Seq1 -> INT_MIN row, plus a normal row
    ("AB-CD*EF", ["-CD", "B-C*"]),   # non-letter bytes
]


@pytest.mark.parametrize("seq1,seqs", EDGE)
def test_edge_shapes(seq1, seqs):
    p = Problem.from_parts((5, 1, 2, 3), seq1, seqs)
    ref = solve_cpu(p, "naive")
    for world in (1, 2):
        assert np.array_equal(plan.execute_blob_cpu(plan.plan_pack(p, world)), ref)


@pytest.mark.parametrize("seed", range(8))
def test_random_vs_oracle(seed):
    rng = np.random.default_rng(100 + seed)
    w = tuple(int(x) for x in rng.integers(0, 30, size=4))
    len1 = int(rng.integers(30, 1400))
    p = random_problem(10, len1=len1, len2_range=(1, min(len1 + 3, 1300)), weights=w, seed=seed)
    assert np.array_equal(plan.execute_blob_cpu(plan.plan_pack(p, 1 + seed % 3)), solve_cpu(p, "fast"))


def test_ties_heavy():
    # constant sequences: every candidate ties; first (n=0,k=0) must win
    p = Problem.from_parts((3, 1, 1, 1), "A" * 700, ["A" * 5, "A" * 300, "B" * 40])
    assert np.array_equal(plan.execute_blob_cpu(plan.plan_pack(p)), solve_cpu(p, "fast"))


def test_long_sequences_beyond_reference_caps():
    p = random_problem(2, len1=5000, len2_range=(2500, 3000), seed=9)
    assert np.array_equal(plan.execute_blob_cpu(plan.plan_pack(p)), solve_cpu(p, "fast"))


def test_unsupported_inputs_are_reported():
    with pytest.raises(NativeError) as e:
        plan.plan_pack(Problem.from_parts((1, 1, 1, 1), "A" * 10000, ["B" * 9000]))
    assert e.value.code == SA_ERR_UNSUPPORTED
    with pytest.raises(NativeError) as e:
        plan.plan_pack(Problem.from_parts((1 << 28, 1, 1, 1), "A" * 100, ["B" * 50]))
    assert e.value.code == SA_ERR_UNSUPPORTED
    with pytest.raises(NativeError) as e:
        plan.plan_pack(Problem.from_parts((1, 1, 1, 1), "A-*+/=", ["B!?"]))  # > 5 distinct non-letters
    assert e.value.code == SA_ERR_UNSUPPORTED


# ---- PAIR mode planning + bit-exact CPU emulation of the u16x2 arithmetic

@pytest.mark.parametrize("i", [1, 2, 3, 4, 5, 6])
def test_pair_mode_walk_matches_golden(shipped_inputs, i):
    path, gold = shipped_inputs[i]
    blob = plan.plan_pack(Problem.from_file(path), 1, mode="pair")
    assert blob.pair_mode and blob.nitems == (Problem.from_file(path).nseq + 1) // 2
    assert format_results(plan.execute_blob_cpu(blob)) == gold.read_bytes()


@pytest.mark.parametrize("seq1,seqs", EDGE)
def test_pair_mode_edge_shapes(seq1, seqs):
    p = Problem.from_parts((5, 1, 2, 3), seq1, seqs)
    assert np.array_equal(plan.execute_blob_cpu(plan.plan_pack(p, 1, mode="pair")), solve_cpu(p, "naive"))


@pytest.mark.parametrize("seed", range(4))
def test_pair_mode_random(seed):
    rng = np.random.default_rng(400 + seed)
    w = tuple(int(x) for x in rng.integers(-5, 200, size=4))  # negative weights shift the bias
    len1 = int(rng.integers(30, 1300))
    p = random_problem(9, len1=len1, len2_range=(1, min(len1 + 2, 1200)), weights=w, seed=seed)
    assert np.array_equal(plan.execute_blob_cpu(plan.plan_pack(p, 1, mode="pair")), solve_cpu(p, "fast"))


def test_pair_mode_auto_threshold_and_limits():
    small = random_problem(50, len1=200, len2_range=(10, 100), seed=1)
    assert not plan.plan_pack(small).pair_mode                      # too few sequences
    assert not plan.plan_pack(random_problem(1200, len1=200, len2_range=(10, 100), seed=1), 2).pair_mode  # multi-GPU
    assert plan.plan_pack(random_problem(1200, len1=200, len2_range=(10, 100), seed=1)).mode in ("pair", "stream")
    with pytest.raises(NativeError) as e:
        plan.plan_pack(random_problem(4, len1=200, len2_range=(10, 100), weights=(70000, 1, 1, 1), seed=1), mode="pair")
    assert e.value.code == SA_ERR_UNSUPPORTED
    with pytest.raises(NativeError):  # profile of a 3000-row pair cannot fit in shared memory
        plan.plan_pack(random_problem(2, len1=4000, len2_range=(3000, 3000), seed=1), mode="pair")


# ---- STREAM mode (warp-autonomous u16x2): planning + emulation

@pytest.mark.parametrize("i", [1, 2, 3, 4, 5, 6])
def test_stream_mode_walk_matches_golden(shipped_inputs, i):
    path, gold = shipped_inputs[i]
    blob = plan.plan_pack(Problem.from_file(path), 1, mode="stream")
    assert blob.stream_mode and blob.header.npairs == (Problem.from_file(path).nseq + 1) // 2
    assert format_results(plan.execute_blob_cpu(blob)) == gold.read_bytes()


@pytest.mark.parametrize("seq1,seqs", EDGE)
def test_stream_mode_edge_shapes(seq1, seqs):
    p = Problem.from_parts((5, 1, 2, 3), seq1, seqs)
    assert np.array_equal(plan.execute_blob_cpu(plan.plan_pack(p, 1, mode="stream")), solve_cpu(p, "naive"))


@pytest.mark.parametrize("seed", range(4))
def test_stream_mode_random(seed):
    rng = np.random.default_rng(500 + seed)
    w = tuple(int(x) for x in rng.integers(-3, 14, size=4))
    len1 = int(rng.integers(30, 1300))
    p = random_problem(9, len1=len1, len2_range=(1, min(len1 + 2, 1200)), weights=w, seed=seed)
    assert np.array_equal(plan.execute_blob_cpu(plan.plan_pack(p, 1, mode="stream")), solve_cpu(p, "fast"))


def test_auto_prefers_stream_then_pair_then_tile():
    big = random_problem(12000, len1=600, len2_range=(10, 100), weights=(10, 2, 3, 4), seed=1)
    assert plan.plan_pack(big).mode == "stream"            # >= 16384 warp items
    assert plan.plan_pack(big, 8).mode == "tile"           # ... but not per GPU on 8 GPUs
    assert plan.plan_pack(big, 8, mode="stream").mode == "stream"
    ok = random_problem(1200, len1=600, len2_range=(10, 500), weights=(10, 2, 3, 4), seed=1)
    assert plan.plan_pack(ok).mode == "pair"               # too few items for warp-autonomous tiles
    assert plan.plan_pack(ok, mode="stream").mode == "stream"
    heavy = random_problem(1200, len1=600, len2_range=(10, 500), weights=(200, 2, 3, 4), seed=1)  # 512 rows x 204 > 32767
    assert plan.plan_pack(heavy).mode == "pair"
    with pytest.raises(NativeError):
        plan.plan_pack(heavy, mode="stream")
    huge = random_problem(1200, len1=600, len2_range=(10, 500), weights=(40000, 2, 3, 4), seed=1)
    assert plan.plan_pack(huge).mode == "tile"


def test_stream_mode_partitions_whole_pairs_over_ranks():
    p = random_problem(1500, len1=700, len2_range=(20, 600), seed=8)
    assert plan.plan_pack(p, 4).mode == "tile"  # AUTO: too few items per GPU
    blob = plan.plan_pack(p, 4, mode="stream")
    assert blob.mode == "stream" and blob.header.world == 4
    h, tab, items = blob.header, blob.seq_table(), blob.items()
    pairs = blob.pair_records()
    loads = []
    for r in range(4):
        sl = items[h.rank_first_item[r] : h.rank_first_item[r + 1]]
        # a rank's pairs are one contiguous slice of the record table, its letters one byte range
        assert sl["seq"].min() == h.rank_first_pair[r] and sl["seq"].max() == h.rank_first_pair[r + 1] - 1
        mine = pairs[h.rank_first_pair[r] : h.rank_first_pair[r + 1]]
        assert mine["off_a"].min() >= h.rank_seq2_lo[r] and (mine["off_a"] + mine["crow"]).max() <= h.rank_seq2_hi[r]
        assert np.all(mine["home"] == r)
        for pid in np.unique(sl["seq"]):
            a, b = int(pairs["a"][pid]), int(pairs["b"][pid])
            assert int(tab["home"][a]) == r and (b == 0xFFFFFFFF or int(tab["home"][b]) == r)
        loads.append(int((tab["crow"][pairs["a"][sl["seq"]]].astype(np.int64) * 256).sum()))
    assert max(loads) <= 1.05 * (sum(loads) / 4)
    assert np.array_equal(plan.execute_blob_cpu(blob), solve_cpu(p, "fast"))
    with pytest.raises(NativeError):
        plan.plan_pack(p, 4, mode="pair")  # PAIR (CTA-per-pair) is single-GPU


def _sparse_rank_walk(blob, world):
    """Score every rank's items from a blob that holds ONLY that rank's ranges (sa_blob_rank_ranges),
    max-reduce the keys, finalise on the full blob: what the fabric kernels / the CPU world do."""
    import ctypes as C

    from mpi_openmp_cuda_b200.ops import _native as N
    from mpi_openmp_cuda_b200.parallel.cpu_world import rank_ranges

    full = blob.numpy()[: blob.nbytes]
    h = blob.header
    keys = np.zeros(max(blob.nseq, 1), dtype=np.uint64)
    moved = []
    for r in range(world):
        sparse = np.zeros(blob.nbytes, dtype=np.uint8)
        sparse[:1024] = full[:1024]
        rr = rank_ranges(h, r)
        for a, b in rr:
            sparse[a:b] = full[a:b]
        moved.append(sum(b - a for a, b in rr))
        mine = np.zeros_like(keys)
        N.check(N.core().sa_blob_score_items_cpu(C.c_void_p(sparse.ctypes.data), blob.nbytes, int(h.rank_first_item[r]),
                                                 int(h.rank_first_item[r + 1]), mine.ctypes.data_as(C.POINTER(C.c_uint64))))
        keys = np.maximum(keys, mine)
    out = np.zeros((max(blob.nseq, 1), 4), dtype=np.int32)
    N.check(N.core().sa_blob_finalize_cpu(C.c_void_p(full.ctypes.data), blob.nbytes, keys.ctypes.data_as(C.POINTER(C.c_uint64)),
                                          out.ctypes.data_as(C.POINTER(C.c_int32))))
    return out[: blob.nseq, :3], moved


@pytest.mark.parametrize("mode", ["tile", "stream"])
def test_each_rank_needs_only_its_own_ranges(mode):
    p = random_problem(600, len1=700, len2_range=(20, 600), seed=11)
    blob = plan.plan_pack(p, 4, mode=mode)
    res, moved = _sparse_rank_walk(blob, 4)
    assert np.array_equal(res, solve_cpu(p, "fast"))
    # a scatter, not a broadcast: what the four ranks receive adds up to about one blob, not four
    assert sum(moved) < 1.3 * blob.nbytes


def test_split_tiles_deals_one_sequence_over_several_ranks(shipped_inputs):
    p = Problem.from_file(shipped_inputs[3][0])
    blob = plan.plan_pack(p, 4, mode="tile", split_tiles=True)
    h, items = blob.header, blob.items()
    owners = {}
    for r in range(4):
        for s in items[h.rank_first_item[r] : h.rank_first_item[r + 1]]["seq"]:
            owners.setdefault(int(s), set()).add(r)
    assert any(len(v) > 1 for v in owners.values()), "no sequence was split over ranks"
    res, _ = _sparse_rank_walk(blob, 4)
    assert np.array_equal(res, solve_cpu(p, "fast"))


@pytest.mark.parametrize("mode", ["tile", "pair", "stream"])
def test_empty_sequence_has_no_candidates_in_every_layer(mode):
    """An empty Seq2 (the parser cannot produce one, from_parts can) is a row without candidates:
    (INT_MIN, 0, 0) from both oracles and from the walk of every mode, as for len2 > len1."""
    seqs = ["", "HELLO", "", "WORLD", "HELLOWORLDHELLO"]
    p = Problem.from_parts((5, 1, 2, 3), "HELLOWORLD", seqs)
    naive, fast = solve_cpu(p, "naive"), solve_cpu(p, "fast")
    assert np.array_equal(naive, fast)
    assert naive[0].tolist() == [-2147483648, 0, 0] and naive[2].tolist() == [-2147483648, 0, 0]
    assert p.candidates() == 5 * 5 * 2
    blob = plan.plan_pack(p, mode=mode)
    assert blob.seq_table()["ntiles"][0] == 0 and blob.seq_table()["ntiles"][2] == 0
    assert np.array_equal(plan.execute_blob_cpu(blob), naive)


@pytest.mark.parametrize("mode", ["tile", "pair", "stream", "auto"])
def test_packing_a_range_of_sequences_scores_exactly_those_rows(shipped_inputs, mode):
    """sa_plan_pack_range: what the native session packs range by range for large batches. Row i of a
    range's launch must be sequence first + i, whatever the range, including ranges with sequences that
    have no candidates and the empty range."""
    p = Problem.from_file(shipped_inputs[3][0])
    want = solve_cpu(p, "fast")
    for first, count in ((0, p.nseq), (0, 7), (5, 11), (p.nseq - 3, 3), (9, 1), (4, 0)):
        if count == 0 and mode in ("pair", "stream"):
            continue  # a forced u16x2 mode needs at least one sequence with candidates
        blob = plan.plan_pack_range(p, first, count, mode=mode)
        assert blob.nseq == count
        got = plan.execute_blob_cpu(blob)
        assert np.array_equal(got, want[first:first + count])
    whole = plan.plan_pack(p, mode=mode)
    ranged = plan.plan_pack_range(p, 0, p.nseq, mode=mode)
    assert bytes(whole.numpy()[: whole.nbytes]) == bytes(ranged.numpy()[: ranged.nbytes])
    with pytest.raises(Exception):
        plan.plan_pack_range(p, p.nseq - 1, 2)
