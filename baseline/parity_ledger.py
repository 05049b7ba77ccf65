#!/usr/bin/env python
"""Per-row parity ledger: where does the UNMODIFIED reference binary disagree with the section 8.2
contract, is the disagreement stable across runs, and does it vanish when the host compiler
zero-fills automatic variables (the analysis-only `final_zeroinit` build, see build_ref.sh)?

    python baseline/parity_ledger.py --inputs 1 3 --repeat 5 --out gpurun_out/parity_ledger.json

Pure Python + baseline/refgen.py; loads nothing native from this repository."""
from __future__ import annotations

import argparse
import json
import re
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent
sys.path.insert(0, str(ROOT))

from baseline import refgen  # noqa: E402
from baseline.run_ref import REF, ensure_built, run_once  # noqa: E402

ROW = re.compile(rb"#(\d+): score: (-?\d+), n: (-?\d+), k: (-?\d+)")


def rows_of(out: bytes):
    return {int(m.group(1)): (int(m.group(2)), int(m.group(3)), int(m.group(4))) for m in ROW.finditer(out)}


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--inputs", type=int, nargs="+", default=[1, 3])
    ap.add_argument("--repeat", type=int, default=5)
    ap.add_argument("--np", type=int, default=1)
    ap.add_argument("--variants", nargs="+", default=["final", "final_zeroinit"])
    ap.add_argument("--out", default=str(ROOT / "gpurun_out" / "parity_ledger.json"))
    a = ap.parse_args()
    why = ensure_built()
    if why:
        print(json.dumps({"unavailable": why}))
        return 0
    ledger = []
    for i in a.inputs:
        data = (ROOT / "data" / f"input{i}.txt").read_bytes()
        w, s1, seqs = refgen.parse_text(data)
        want = refgen.solve(w, s1, seqs)
        per_variant = {}
        for v in a.variants:
            if not (REF / "bin" / v).exists():
                continue
            runs = []
            for _ in range(a.repeat):
                dt, rc, out, err, dev = run_once(data, a.np, v, timeout=1800, probe=False)
                runs.append(rows_of(out) if rc == 0 else None)
            per_variant[v] = runs
        for row, exp in enumerate(want):
            entry = {"input": i, "row": row, "len2": len(seqs[row]), "contract": list(exp)}
            differs = False
            for v, runs in per_variant.items():
                seen = [list(r.get(row, ())) if r is not None else None for r in runs]
                distinct = [list(x) for x in sorted({tuple(s) for s in seen if s})]
                entry[v] = {"distinct_outputs": distinct, "stable": len(distinct) == 1,
                            "matches_contract": distinct == [list(exp)]}
                differs |= distinct != [list(exp)]
            if differs:
                ledger.append(entry)
        print(f"input{i}: {sum(1 for e in ledger if e['input'] == i)} rows differ in some variant", flush=True)
    Path(a.out).parent.mkdir(parents=True, exist_ok=True)
    Path(a.out).write_text(json.dumps({"np": a.np, "repeat": a.repeat, "rows": ledger}, indent=1))
    return 0


if __name__ == "__main__":
    sys.exit(main())
