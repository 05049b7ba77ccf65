"""Pair classes (reference: group lists main.c:59-60, table build main.c:14-44)."""
import itertools
import string

from mpi_openmp_cuda_b200.ops import table


def test_all_letter_pairs_match_group_scan():
    for a, b in itertools.product(string.ascii_uppercase, repeat=2):
        assert table.pair_class(a, b) == table.pair_class_by_scan(a, b), (a, b)


def test_symmetry_and_identity():
    for a, b in itertools.product(string.ascii_uppercase, repeat=2):
        assert table.pair_class(a, b) == table.pair_class(b, a)
    for a in string.ascii_uppercase:
        assert table.pair_class(a, a) == 0


def test_known_pairs():
    assert table.pair_class("N", "D") == 1  # NDEQ
    assert table.pair_class("S", "G") == 2  # SAG
    assert table.pair_class("S", "A") == 1  # STA wins over SAG/CSA (type-1 checked first)
    assert table.pair_class("B", "A") == 3
    assert table.pair_class("W", "Y") == 1  # FYW


def test_non_letters():
    assert table.pair_class("-", "-") == 0
    assert table.pair_class("-", "A") == 3
    assert table.pair_class("*", "-") == 3


def test_value_table_signs():
    t = table.value_table((10, 2, 3, 4))
    assert t[0, 0] == 10
    assert t[ord("N") - 65, ord("D") - 65] == -2
    assert t[ord("S") - 65, ord("G") - 65] == -3
    assert t[ord("B") - 65, ord("A") - 65] == -4
