"""Host-side sanitizer run: the whole core library under ASan + UBSan through
tests/native/core_selftest.cpp in quick mode. The full runs (`make asan`, `make tsan`, several
minutes) are recorded in profiles/sanitizers.txt."""
import os
import shutil
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
SRC = ["tests/native/core_selftest.cpp", "csrc/core/problem.cpp", "csrc/core/oracle.cpp", "csrc/core/plan.cpp"]


def test_core_library_is_clean_under_asan_ubsan(tmp_path):
    cxx = "/usr/bin/g++" if Path("/usr/bin/g++").exists() else shutil.which("g++")
    exe = tmp_path / "selftest_asan"
    build = subprocess.run([cxx, "-O1", "-g", "-std=c++17", "-Iinclude", "-pthread", "-fsanitize=address,undefined",
                            "-fno-sanitize-recover=undefined", "-o", str(exe), *SRC], cwd=ROOT, capture_output=True, text=True)
    if build.returncode != 0 and ("-lasan" in build.stderr or "libasan" in build.stderr):
        pytest.skip("this compiler ships no sanitizer runtime")
    assert build.returncode == 0, build.stderr[-2000:]
    run = subprocess.run([str(exe), "data/input1.txt", "data/input5.txt", "data/input6.txt"], cwd=ROOT, capture_output=True,
                         text=True, timeout=900, env=dict(os.environ, SA_SELFTEST_QUICK="1", ASAN_OPTIONS="detect_leaks=1"))
    assert run.returncode == 0, (run.stdout + run.stderr)[-3000:]
    assert "core_selftest: ok" in run.stdout
