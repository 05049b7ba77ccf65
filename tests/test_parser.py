"""Input grammar (reference: main.c:76-108): whitespace tokens, CRLF, double blanks, case."""
import pytest

from mpi_openmp_cuda_b200 import Problem
from mpi_openmp_cuda_b200.ops._native import NativeError


def test_basic_tokens():
    p = Problem.parse(b"10 2 3 4\nAPQRSBATAV\n2\nASQ\nrstb\n")
    assert p.weights == (10, 2, 3, 4)
    assert p.seq1 == b"APQRSBATAV"
    assert p.nseq == 2
    assert p.seq2(0) == b"ASQ"
    assert p.seq2(1) == b"RSTB"  # a-z upper-cased (main.c:102-106)


def test_crlf_double_blanks_no_final_newline():
    p = Problem.parse(b"2  2  1  10\r\nabcDEF\r\n1\r\nxyZ")
    assert p.weights == (2, 2, 1, 10)
    assert p.seq1 == b"ABCDEF"
    assert p.seq2(0) == b"XYZ"


def test_tokens_on_one_line_and_tabs():
    p = Problem.parse(b"1 2 3 4 HELLO 2 HE\tLLO")
    assert p.seq2(0) == b"HE" and p.seq2(1) == b"LLO"


def test_negative_and_signed_weights():
    p = Problem.parse(b"-1 +2 3 4 AB 0")
    assert p.weights == (-1, 2, 3, 4)
    assert p.nseq == 0


def test_non_letters_pass_through():
    p = Problem.parse(b"1 1 1 1 AB-C*D 1 a-b")
    assert p.seq1 == b"AB-C*D"
    assert p.seq2(0) == b"A-B"


@pytest.mark.parametrize("text", [b"", b"1 2 3", b"1 2 3 x", b"1 2 3 4", b"1 2 3 4 ABC", b"1 2 3 4 ABC 2 AB", b"1 2 3 4 ABC -1"])
def test_missing_tokens_are_errors(text):
    with pytest.raises(NativeError):
        Problem.parse(text)


def test_shipped_inputs_parse(shipped_inputs):
    expect = {1: (51, 10), 2: (344, 1), 3: (1489, 32), 4: (2976, 30), 5: (10, 1), 6: (26, 5)}
    cands = {1: 5006, 2: 343, 3: 10733894, 4: 4457709, 5: 9, 6: 690}
    for i, (path, _) in shipped_inputs.items():
        p = Problem.from_file(path)
        assert (len(p.seq1), p.nseq) == expect[i]
        assert p.candidates() == cands[i]
        assert p.seq1 == p.seq1.upper()


def test_serialize_roundtrip():
    p = Problem.from_parts((4, 3, 2, 10), "abcXYZ", ["ab", "XyZ"])
    q = Problem.parse(p.serialize())
    assert q.weights == p.weights and q.seq1 == b"ABCXYZ" and q.seq2(1) == b"XYZ"


def test_hostile_sequence_count_is_a_parse_error_not_an_allocation():
    # found by `make asan`: the count must not size an allocation before the tokens are seen
    with pytest.raises(NativeError):
        Problem.parse(b"1 1 1 1 A 2147483647 A")
