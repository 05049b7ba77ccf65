"""`python -m mpi_openmp_cuda_b200 [--cpu] [--gpus N] < input.txt` — the `./final` contract from Python:
the reference's stdin grammar in, `#i: score: S, n: N, k: K` lines out (main.c:46-244)."""
from __future__ import annotations

import argparse
import sys


def main() -> int:
    ap = argparse.ArgumentParser(prog="python -m mpi_openmp_cuda_b200")
    ap.add_argument("--cpu", action="store_true", help="CPU oracle instead of the GPU")
    ap.add_argument("--gpus", type=int, default=1, help="number of GPUs (one process per GPU when > 1)")
    a = ap.parse_args()
    if a.gpus > 1:
        from .parallel import launcher

        sys.argv = [sys.argv[0], "--gpus", str(a.gpus)] + (["--cpu"] if a.cpu else [])
        return launcher.main()
    from . import format_results, parse_text, solve

    try:
        problem = parse_text(sys.stdin.buffer.read())
    except Exception as e:  # malformed input: message on stderr, non-zero exit, nothing on stdout
        print(f"error: {e}", file=sys.stderr)
        return 1
    sys.stdout.buffer.write(format_results(solve(problem, device="cpu" if a.cpu else None)))
    return 0


if __name__ == "__main__":
    sys.exit(main())
