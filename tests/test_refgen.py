"""The reference arm's pure-Python helpers (baseline/refgen.py): same input bytes as the native
generator, same answers as the goldens, and no native library of this repository loaded."""
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

from baseline import refgen  # noqa: E402


@pytest.mark.parametrize("i", [1, 2, 3, 4, 5, 6])
def test_numpy_oracle_matches_golden(i):
    data = (ROOT / "data" / f"input{i}.txt").read_bytes()
    out = refgen.format_results(refgen.solve(*refgen.parse_text(data)))
    assert out == (ROOT / "data" / "golden" / f"input{i}.out").read_bytes()


def test_generator_is_byte_identical_to_the_native_one():
    from mpi_openmp_cuda_b200 import Problem
    from mpi_openmp_cuda_b200.models.synthetic import scaled_input3

    txt = refgen.scaled_text(ROOT / "data" / "input3.txt", 3, seed=1)
    p = scaled_input3(3, ROOT / "data" / "input3.txt", seed=1)
    assert p.serialize() == txt
    w, s1, seqs = refgen.parse_text(txt)
    assert refgen.candidates(s1, seqs) == p.candidates() == Problem.parse(txt).candidates()


def test_edge_cases_match_the_native_oracle():
    import numpy as np

    from mpi_openmp_cuda_b200 import Problem
    from mpi_openmp_cuda_b200.ops.oracle import solve_cpu

    seqs = [b"HELLOWORLD", b"OWRL", b"HELLOWORLDX", b"A", b"DL", b"HELLOWORL"]
    p = Problem.from_parts((10, 2, 3, 4), b"HELLOWORLD", seqs)
    want = solve_cpu(p, "naive")
    got = np.array(refgen.solve((10, 2, 3, 4), b"HELLOWORLD", seqs), dtype=np.int64)
    assert np.array_equal(got, want)


def test_reference_arm_loads_no_native_code_of_this_repo():
    code = (
        "import sys; sys.argv=['bench.py','--impl','reference','--steps','1','--warmup','3'];"
        "sys.path.insert(0, %r); import bench; from baseline import refgen, run_ref, contract;"
        "a=bench.parse_args(); d=refgen.scaled_text(%r, 2, seed=1); run_ref.expected_output(d);"
        "print([l.split()[-1] for l in open('/proc/self/maps') if 'seqalign' in l])"
    ) % (str(ROOT), str(ROOT / "data" / "input1.txt"))
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    assert out.stdout.strip().splitlines()[-1] == "[]"
