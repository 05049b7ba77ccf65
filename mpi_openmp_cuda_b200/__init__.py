"""B200-native exhaustive sequence-alignment solver (capabilities of nmiz1987/MPI-OPENMP-CUDA).

Layout: `utils` (grammar, printer, build, timing), `ops` (pair table, oracle, planner,
sm_100a scoring kernel bindings), `models` (the solver object + synthetic workloads),
`parallel` (one-process-per-GPU fabric: in-kernel broadcast and argmax reduce).
"""
from .utils.parser import Problem, parse_text
from .utils.format import format_results

__all__ = ["Problem", "parse_text", "format_results", "solve", "solve_text"]


def solve(problem, device: str | None = None):
    """Solve on the GPU when one is present (else the CPU oracle) -> int32[nseq,3]."""
    from .models.alignment import solve as _solve

    return _solve(problem, device=device)


def solve_text(text, device: str | None = None) -> bytes:
    """stdin bytes -> stdout bytes, the reference's whole contract (main.c:46-244)."""
    p = parse_text(text)
    return format_results(solve(p, device=device))
