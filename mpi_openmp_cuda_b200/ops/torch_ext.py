"""`torch.ops.seqalign.*` — the PyTorch extension binding (csrc/torch/binding.cpp).

Built in-tree with torch.utils.cpp_extension (no JIT cache outside the repo) and linked against
libseqalign_core.so / libseqalign_cuda.so. `load()` returns the ops namespace."""
from __future__ import annotations

import os
from functools import lru_cache

from ..utils import build as _build

EXT_DIR = _build.LIB_DIR / "torch_ext"


def build(verbose: bool = False):
    import torch.utils.cpp_extension as ce

    _build.ensure_core()
    _build.ensure_cuda()
    EXT_DIR.mkdir(parents=True, exist_ok=True)
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0a")
    return ce.load(
        name="seqalign_torch",
        sources=[str(_build.REPO_ROOT / "csrc" / "torch" / "binding.cpp")],
        extra_include_paths=[str(_build.REPO_ROOT / "include")],
        extra_cflags=["-O2", "-std=c++17"],
        extra_ldflags=[f"-L{_build.LIB_DIR}", "-lseqalign_cuda", "-lseqalign_core", f"-Wl,-rpath,{_build.LIB_DIR}",
                       "-Wl,-rpath,$ORIGIN/.."],
        build_directory=str(EXT_DIR),
        with_cuda=True,
        verbose=verbose,
    )


@lru_cache(maxsize=None)
def load():
    """Import the built extension (building it when missing) and return `torch.ops.seqalign`."""
    import torch

    so = EXT_DIR / "seqalign_torch.so"
    if so.exists():
        from . import _native

        _native.core()
        torch.ops.load_library(str(so))
    else:
        build()
    return torch.ops.seqalign
