"""The MPI stand-in used ONLY to build/run the unmodified reference (the image has no MPI).
Multi-process CPU test: real processes, the reference's collective sequence, stdin to rank 0."""
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
SHIM = ROOT / "baseline" / "mpi_shim"


@pytest.fixture(scope="module")
def shim_build(tmp_path_factory):
    d = tmp_path_factory.mktemp("shim")
    subprocess.run(["g++", "-O2", "-fPIC", "-shared", "-o", str(d / "libmpishim.so"), str(SHIM / "mpi_shim.cpp"),
                    f"-I{SHIM}", "-lpthread"], check=True)
    subprocess.run(["g++", "-O2", "-o", str(d / "mpiexec"), str(SHIM / "mpiexec.cpp"), f"-I{SHIM}", "-lpthread"], check=True)
    subprocess.run(["g++", "-O2", "-o", str(d / "selftest"), str(SHIM / "selftest.cpp"), f"-I{SHIM}", f"-L{d}",
                    "-lmpishim", "-lpthread", f"-Wl,-rpath,{d}"], check=True)
    return d


@pytest.mark.parametrize("np_", [1, 2, 4, 8])
def test_collectives_across_processes(shim_build, np_):
    out = subprocess.run([str(shim_build / "mpiexec"), "-np", str(np_), str(shim_build / "selftest")],
                         input=b"hello from stdin\n", capture_output=True, timeout=60)
    assert out.returncode == 0, out.stderr
    text = out.stdout.decode()
    assert f"world {np_}: PASS" in text
    assert "rank0 stdin: hello from stdin" in text
    assert "unexpectedly" not in text


def test_dead_rank_does_not_hang_the_job(shim_build, tmp_path):
    prog = tmp_path / "die.sh"
    prog.write_text("#!/bin/bash\nif [ \"$SHIM_RANK\" = 1 ]; then exit 3; fi\nsleep 30\n")
    prog.chmod(0o755)
    out = subprocess.run([str(shim_build / "mpiexec"), "-np", "3", str(prog)], capture_output=True, timeout=20)
    assert out.returncode == 3
