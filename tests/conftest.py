import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
DATA = ROOT / "data"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch

        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session", autouse=True)
def _native_built():
    """CPU library is built on demand; the CUDA library must already be in-tree on a GPU box."""
    from mpi_openmp_cuda_b200.utils import build

    build.ensure_core()
    if not (ROOT / "final").exists():
        build.build_all(("core", "final")) if build.lib_path("seqalign_cuda").exists() else None
    yield


@pytest.fixture(scope="session")
def shipped_inputs():
    return {i: (DATA / f"input{i}.txt", DATA / "golden" / f"input{i}.out") for i in range(1, 7)}
