"""Property tests (hypothesis): every execution mode's CPU walk equals the literal oracle on
arbitrary small problems — lengths 1..3, equal lengths, longer-than-Seq1, ties, lower case,
non-letter bytes, zero and negative weights all fall out of the generator."""
import numpy as np
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from mpi_openmp_cuda_b200 import Problem
from mpi_openmp_cuda_b200.ops import plan
from mpi_openmp_cuda_b200.ops._native import NativeError, SA_ERR_UNSUPPORTED
from mpi_openmp_cuda_b200.ops.oracle import solve_cpu

letters = st.text(alphabet="ABCDEFGHIKLMNPQRSTVWY" + "xz-", min_size=1, max_size=40)
weights = st.tuples(*[st.integers(min_value=-3, max_value=25)] * 4)


@settings(max_examples=120, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(w=weights, seq1=st.text(alphabet="ABCDEFGHIKLMNPQRSTVWY-", min_size=1, max_size=60), seqs=st.lists(letters, min_size=1, max_size=6),
       world=st.integers(min_value=1, max_value=4))
def test_every_mode_matches_the_literal_oracle(w, seq1, seqs, world):
    p = Problem.from_parts(w, seq1, seqs)
    ref = solve_cpu(p, "naive")
    assert np.array_equal(solve_cpu(p, "fast"), ref)
    for mode in ("tile", "pair", "stream"):
        if mode == "pair" and world != 1:
            continue
        try:
            blob = plan.plan_pack(p, world, mode=mode)
        except NativeError as e:
            assert e.code == SA_ERR_UNSUPPORTED  # e.g. no sequence fits Seq1: nothing to pair
            continue
        assert np.array_equal(plan.execute_blob_cpu(blob), ref), mode


@settings(max_examples=60, deadline=None)
@given(w=weights, seq1=st.text(alphabet="ACDEFGHIKLMNPQRSTVWY", min_size=2, max_size=50), seq2=st.text(alphabet="ACDEFGHIKLMNPQRSTVWY", min_size=1, max_size=50))
def test_score_formula_on_the_reported_candidate(w, seq1, seq2):
    """The reported (n, k) really scores what is reported (recomputed from the pair classes)."""
    from mpi_openmp_cuda_b200.ops.table import pair_class

    p = Problem.from_parts(w, seq1, [seq2])
    score, n, k = (int(x) for x in solve_cpu(p, "fast")[0])
    if len(seq2) > len(seq1):
        assert (score, n, k) == (-2147483648, 0, 0)
        return
    vals = (w[0], -w[1], -w[2], -w[3])
    total = 0
    for i, c in enumerate(seq2):
        j = i + n if (k == 0 or i < k) else i + n + 1
        total += vals[pair_class(c, seq1[j])]
    assert total == score


@settings(max_examples=80, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(w=weights, seq1=st.text(alphabet="ABCDEFGHIKLMNPQRSTVWY-", min_size=1, max_size=300), seqs=st.lists(letters, min_size=1, max_size=9),
       cut=st.tuples(st.integers(min_value=0, max_value=9), st.integers(min_value=0, max_value=9)), split=st.booleans(),
       world=st.integers(min_value=1, max_value=3))
def test_ranges_narrow_tiles_and_split_tiles_match_the_oracle(w, seq1, seqs, cut, split, world):
    """Round-2 planner paths: a range of the sequences packed on its own (what the session does for large
    batches), the latency mode's narrow tiles (small problems always get them), tiles of one sequence dealt
    over several ranks. Whatever the cut, row i of a range is sequence first + i of the oracle."""
    p = Problem.from_parts(w, seq1, seqs)
    ref = solve_cpu(p, "fast")
    first = min(cut[0], p.nseq)
    count = min(cut[1], p.nseq - first)
    blob = plan.plan_pack_range(p, first, count, mode="tile")
    assert np.array_equal(plan.execute_blob_cpu(blob), ref[first:first + count])
    whole = plan.plan_pack(p, world, mode="tile", split_tiles=split)
    assert np.array_equal(plan.execute_blob_cpu(whole), ref)
