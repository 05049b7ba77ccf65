"""sm_100a scoring kernel vs the CPU oracle (plain reference of the same op) and the goldens."""
import ctypes as C

import numpy as np
import pytest

from mpi_openmp_cuda_b200 import Problem, format_results, solve_text
from mpi_openmp_cuda_b200.models.synthetic import random_problem
from mpi_openmp_cuda_b200.ops import plan
from mpi_openmp_cuda_b200.ops.oracle import solve_cpu

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def solver():
    from mpi_openmp_cuda_b200.models.alignment import AlignmentSolver

    return AlignmentSolver("cuda:0")


def test_cuda_library_is_loaded_not_a_fallback():
    from mpi_openmp_cuda_b200.ops import _native

    lib = _native.cuda()
    assert "libseqalign_cuda.so" in lib._name


@pytest.mark.parametrize("i", [1, 2, 3, 4, 5, 6])
def test_golden_outputs(solver, shipped_inputs, i):
    path, gold = shipped_inputs[i]
    assert format_results(solver.solve(Problem.from_file(path))) == gold.read_bytes()


def test_solve_text_contract(shipped_inputs):
    path, gold = shipped_inputs[6]
    assert solve_text(path.read_bytes(), device="cuda:0") == gold.read_bytes()


def test_repeated_launches_reuse_self_cleaning_scratch(solver, shipped_inputs):
    from mpi_openmp_cuda_b200.ops.score import launch_count

    p3, p4 = Problem.from_file(shipped_inputs[3][0]), Problem.from_file(shipped_inputs[4][0])
    r3, r4 = solve_cpu(p3), solve_cpu(p4)
    before = launch_count()
    for _ in range(5):
        assert np.array_equal(solver.solve(p3), r3)
        assert np.array_equal(solver.solve(p4), r4)
    assert launch_count() - before == 10  # one kernel per solve


EDGE = [
    ("ABCDEFGH", ["A"]),
    ("ABCDEFGH", ["AB", "ABC"]),
    ("ABCDEFGH", ["ABCDEFGH"]),
    ("ABCDEFGH", ["ABCDEFG"]),
    ("ABC", ["ABCD", "AB"]),
    ("AB-CD*EF", ["-CD", "B-C*"]),
    ("A" * 700, ["A" * 5, "A" * 300, "B" * 40]),
]


@pytest.mark.parametrize("seq1,seqs", EDGE)
def test_edge_shapes(solver, seq1, seqs):
    p = Problem.from_parts((5, 1, 2, 3), seq1, seqs)
    assert np.array_equal(solver.solve(p), solve_cpu(p, "fast"))


@pytest.mark.parametrize("seed", range(6))
def test_random_vs_oracle(solver, seed):
    rng = np.random.default_rng(200 + seed)
    w = tuple(int(x) for x in rng.integers(0, 40, size=4))
    len1 = int(rng.integers(30, 3000))
    p = random_problem(48, len1=len1, len2_range=(1, min(len1 + 3, 2000)), weights=w, seed=seed)
    assert np.array_equal(solver.solve(p), solve_cpu(p, "fast"))


def test_len2_above_1024_is_deterministic(solver):
    # the reference races across two blocks for len2 > 1024 (cudaFunctions.cu:116-168)
    p = random_problem(8, len1=2900, len2_range=(1025, 1999), seed=5)
    ref = solve_cpu(p, "fast")
    for _ in range(3):
        assert np.array_equal(solver.solve(p), ref)


def test_many_sequences(solver):
    p = random_problem(3000, len1=800, len2_range=(20, 500), seed=11)
    assert np.array_equal(solver.solve(p), solve_cpu(p, "fast"))


def test_beyond_reference_caps(solver):
    p = random_problem(3, len1=9000, len2_range=(3000, 6000), seed=2)
    assert np.array_equal(solver.solve(p), solve_cpu(p, "fast"))


def test_c_abi_convenience_entry(shipped_inputs):
    from mpi_openmp_cuda_b200.ops import _native as N

    p = Problem.from_file(shipped_inputs[1][0])
    out = np.zeros((p.nseq, 3), dtype=np.int32)
    N.check(N.cuda().sa_cuda_solve_problem(p.handle, out.ctypes.data_as(C.POINTER(C.c_int32)), 0))
    assert np.array_equal(out, solve_cpu(p, "fast"))


def test_native_session_and_torch_buffer_paths_agree(solver, shipped_inputs):
    for i in (1, 3, 4):
        p = Problem.from_file(shipped_inputs[i][0])
        assert np.array_equal(solver.solve(p), solver.solve_via_torch(p))
    h2d, d2h = solver.session.last_bytes()
    assert h2d > 0 and d2h == 16 * 30


def test_cpu_walk_and_gpu_agree_on_blob(solver, shipped_inputs):
    p = Problem.from_file(shipped_inputs[3][0])
    blob = solver.pack(p)
    assert np.array_equal(solver.solve_packed(blob), plan.execute_blob_cpu(blob))


# ---- PAIR mode (u16x2 lanes, per-pair profile): forced on small inputs, automatic on large batches

@pytest.mark.parametrize("i", [1, 2, 3, 4, 5, 6])
def test_pair_mode_golden_outputs(solver, shipped_inputs, i):
    path, gold = shipped_inputs[i]
    blob = solver.pack(Problem.from_file(path), mode="pair")
    assert blob.pair_mode
    assert format_results(solver.solve_packed(blob)) == gold.read_bytes()


@pytest.mark.parametrize("seq1,seqs", EDGE)
def test_pair_mode_edge_shapes(solver, seq1, seqs):
    p = Problem.from_parts((5, 1, 2, 3), seq1, seqs)
    assert np.array_equal(solver.solve_packed(solver.pack(p, mode="pair")), solve_cpu(p, "fast"))


@pytest.mark.parametrize("seed", range(4))
def test_pair_mode_random_vs_oracle(solver, seed):
    rng = np.random.default_rng(300 + seed)
    w = tuple(int(x) for x in rng.integers(0, 60, size=4))
    len1 = int(rng.integers(30, 2500))
    p = random_problem(61, len1=len1, len2_range=(1, min(len1 + 3, 1200)), weights=w, seed=seed)
    blob = solver.pack(p, mode="pair")
    assert np.array_equal(solver.solve_packed(blob), solve_cpu(p, "fast"))
    assert np.array_equal(plan.execute_blob_cpu(blob), solve_cpu(p, "fast"))


def test_auto_mode_picks_pairs_for_large_batches(solver):
    p = random_problem(2500, len1=1400, len2_range=(40, 1100), seed=31)
    ref = solve_cpu(p, "fast")
    blob = solver.pack(p)
    assert blob.mode == "pair"
    assert np.array_equal(solver.solve_packed(blob), ref)
    assert np.array_equal(solver.solve(p), ref)  # native session, same auto choice
    for mode in ("tile", "stream"):
        b = solver.pack(p, mode=mode)
        assert b.mode == mode
        assert np.array_equal(solver.solve_packed(b), ref)


def test_auto_mode_picks_stream_for_very_large_batches(solver):
    p = random_problem(12000, len1=700, len2_range=(10, 120), seed=32)
    blob = solver.pack(p)
    assert blob.mode == "stream"
    ref = solve_cpu(p, "fast")
    assert np.array_equal(solver.solve_packed(blob), ref)
    assert np.array_equal(solver.solve(p), ref)


def test_large_batches_are_packed_and_launched_range_by_range(solver):
    """The native session overlaps host packing with copy + kernel for batches of >= 16384 sequences:
    several launches, same rows as one launch and as the oracle; buffers survive a change of batch size."""
    p = random_problem(20000, len1=400, len2_range=(1, 90), seed=5)
    ref = solve_cpu(p, "fast")
    assert np.array_equal(solver.solve(p), ref)
    assert solver.session.last_launches() >= 2
    solver.session.set_pipeline(0)
    try:
        assert np.array_equal(solver.solve(p), ref)
        assert solver.session.last_launches() == 1
    finally:
        solver.session.set_pipeline(16384)
    solver.session.set_pipeline(64)  # force the range path on small batches too, with ranges of uneven size
    try:
        for n in (64, 203, 1000, 131):
            q = random_problem(n, len1=300, len2_range=(1, 320), seed=n)
            assert np.array_equal(solver.solve(q), solve_cpu(q, "fast"))
            assert solver.session.last_launches() >= 2
    finally:
        solver.session.set_pipeline(16384)
    assert np.array_equal(solver.solve(p), ref)


def test_pair_mode_limits_fall_back_to_tiles(solver):
    from mpi_openmp_cuda_b200.ops._native import NativeError

    big_w = random_problem(1100, len1=300, len2_range=(20, 200), weights=(40000, 1, 1, 1), seed=5)
    with pytest.raises(NativeError):
        solver.pack(big_w, mode="pair")
    blob = solver.pack(big_w)  # auto: weights do not fit 16 bits -> tiles
    assert not blob.pair_mode
    assert np.array_equal(solver.solve_packed(blob), solve_cpu(big_w, "fast"))


# ---- STREAM mode (warp-autonomous u16x2, private profile blocks, no block barriers)

@pytest.mark.parametrize("i", [1, 2, 3, 4, 5, 6])
def test_stream_mode_golden_outputs(solver, shipped_inputs, i):
    path, gold = shipped_inputs[i]
    blob = solver.pack(Problem.from_file(path), mode="stream")
    assert blob.stream_mode
    for _ in range(2):  # twice: the scratch must come back zeroed
        assert format_results(solver.solve_packed(blob)) == gold.read_bytes()


@pytest.mark.parametrize("seq1,seqs", EDGE)
def test_stream_mode_edge_shapes(solver, seq1, seqs):
    p = Problem.from_parts((5, 1, 2, 3), seq1, seqs)
    assert np.array_equal(solver.solve_packed(solver.pack(p, mode="stream")), solve_cpu(p, "fast"))


@pytest.mark.parametrize("seed", range(4))
def test_stream_mode_random_vs_oracle(solver, seed):
    rng = np.random.default_rng(600 + seed)
    w = tuple(int(x) for x in rng.integers(0, 14, size=4))
    len1 = int(rng.integers(30, 2500))
    p = random_problem(301, len1=len1, len2_range=(1, min(len1 + 3, 1200)), weights=w, seed=seed)
    blob = solver.pack(p, mode="stream")
    assert np.array_equal(solver.solve_packed(blob), solve_cpu(p, "fast"))


# ---- PyTorch extension binding (torch.ops.seqalign.*)

def test_torch_ops_binding(solver, shipped_inputs):
    import torch

    from mpi_openmp_cuda_b200.ops import torch_ext

    ops = torch_ext.load()
    p = Problem.from_file(shipped_inputs[3][0])
    blob = plan.plan_pack(p)
    d_blob = torch.from_numpy(blob.numpy()[: blob.nbytes].copy()).cuda()
    header = torch.from_numpy(blob.numpy()[:1024].copy())
    scratch = torch.zeros(int(ops.scratch_bytes(blob.nseq)), dtype=torch.uint8, device="cuda")
    for _ in range(2):
        out = ops.solve_blob(d_blob, header, scratch)
        assert np.array_equal(out.cpu().numpy()[:, :3], solve_cpu(p, "fast"))
    text = shipped_inputs[1][0].read_bytes().decode("latin-1")
    assert np.array_equal(ops.solve_text(text, 0).numpy(), solve_cpu(Problem.from_file(shipped_inputs[1][0]), "fast"))


def test_degenerate_problems(solver):
    empty = Problem.parse(b"1 2 3 4 AB 0")
    assert solver.solve(empty).shape == (0, 3)
    none_fit = Problem.from_parts((1, 2, 3, 4), "AB", ["ABC", "ABCD"])  # every sequence longer than Seq1
    assert solver.solve(none_fit).tolist() == [[-2147483648, 0, 0]] * 2
    assert solver.solve_packed(solver.pack(none_fit)).tolist() == [[-2147483648, 0, 0]] * 2


@pytest.mark.parametrize("mode", ["tile", "pair", "stream"])
def test_every_kernel_against_the_literal_restatement(solver, shipped_inputs, mode):
    """The fast oracle shares the kernels' prefix-difference algebra; the naive one re-scores every (n, k)
    candidate letter by letter as the reference kernel means to (cudaFunctions.cu:116-168). Small inputs, since
    it is O(len1 * len2^2): the shipped ones that are small, ties, and random problems with every weight sign."""
    problems = [Problem.from_file(shipped_inputs[i][0]) for i in (1, 2, 5, 6)]
    problems.append(Problem.from_parts((3, 1, 2, 5), "ACDEFGHIKL" * 12, ["ACDEFGHIKL" * 3, "GHIKLACDEF" * 2, "L", "AC"]))
    for seed in range(4):
        problems.append(random_problem(24, len1=150, len2_range=(1, 150), seed=100 + seed))
    problems.append(random_problem(12, len1=90, len2_range=(1, 60), seed=7, weights=(0, 5, 0, 1)))
    for p in problems:
        want = solve_cpu(p, "naive")
        blob = solver.pack(p, mode=mode)
        assert blob.mode == mode
        assert np.array_equal(solver.solve_packed(blob), want)
