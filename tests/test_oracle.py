"""CPU oracle vs the problem statement's worked examples and the golden outputs
(SURVEY.md §4 sources of truth 1 and 2)."""
import numpy as np
import pytest

from mpi_openmp_cuda_b200 import Problem, format_results
from mpi_openmp_cuda_b200.models.synthetic import random_problem
from mpi_openmp_cuda_b200.ops.oracle import solve_cpu


def test_pdf_equal_length_example():
    p = Problem.from_parts((10, 2, 3, 4), "APQRSBATAV", ["ASQRSEAVSL"])
    for m in ("naive", "fast"):
        assert solve_cpu(p, m).tolist() == [[36, 0, 0]]


def test_pdf_mutant_example():
    p = Problem.from_parts((10, 2, 3, 4), "HELLOWORLD", ["OWRL"])
    for m in ("naive", "fast"):
        r = solve_cpu(p, m)
        assert r[0, 1] == 4 and r[0, 2] == 2 and r[0, 0] == 40


def test_longer_than_seq1_has_no_candidates():
    p = Problem.from_parts((1, 1, 1, 1), "ABC", ["ABCD"])
    assert solve_cpu(p, "fast").tolist() == [[-2147483648, 0, 0]]
    assert solve_cpu(p, "naive").tolist() == [[-2147483648, 0, 0]]


def test_offset_len1_minus_len2_is_never_tried():
    # the reference's loop is n < len1-len2 (cudaFunctions.cu:116): perfect match at the last offset is missed
    p = Problem.from_parts((10, 1, 1, 1), "XXXXABC", ["ABC"])
    r = solve_cpu(p, "naive")[0]
    assert r[1] < 4


def test_first_maximum_wins_ties():
    # input6-style tie: identical score at many (n,k); smallest n then smallest k, k=0 first
    p = Problem.from_parts((4, 3, 2, 10), "ABCDEFGHIJKLMNOPQRSTUVWXYZ", ["ABCDEF"])
    assert solve_cpu(p, "naive").tolist() == [[24, 0, 0]]


@pytest.mark.parametrize("i", [1, 2, 3, 4, 5, 6])
def test_golden_fast(shipped_inputs, i):
    path, gold = shipped_inputs[i]
    assert format_results(solve_cpu(Problem.from_file(path), "fast")) == gold.read_bytes()


@pytest.mark.parametrize("i", [1, 2, 5, 6])
def test_golden_naive(shipped_inputs, i):
    path, gold = shipped_inputs[i]
    assert format_results(solve_cpu(Problem.from_file(path), "naive")) == gold.read_bytes()


@pytest.mark.parametrize("seed", range(6))
def test_naive_equals_fast_random(seed):
    rng = np.random.default_rng(seed)
    w = tuple(int(x) for x in rng.integers(0, 12, size=4))
    p = random_problem(12, len1=int(rng.integers(20, 90)), len2_range=(1, 60), weights=w, seed=seed)
    assert np.array_equal(solve_cpu(p, "naive"), solve_cpu(p, "fast"))


def test_threads_do_not_change_results(shipped_inputs):
    p = Problem.from_file(shipped_inputs[4][0])
    assert np.array_equal(solve_cpu(p, "fast", threads=1), solve_cpu(p, "fast", threads=8))
