"""bench.py's driver contract on a box without a GPU: one JSON line on stdout whatever happens, the reference arm
reports `unavailable` instead of failing, and both arms quote the same metric and unit (baseline/contract.py)."""
import json
import subprocess
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parents[1]


def last_json(cmd):
    out = subprocess.run([sys.executable, str(ROOT / "bench.py"), *cmd], capture_output=True, cwd=ROOT, timeout=600)
    lines = [l for l in out.stdout.decode().splitlines() if l.strip()]
    assert lines, out.stderr.decode()[-2000:]
    return out.returncode, json.loads(lines[-1])


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_own_arm_without_a_gpu_says_so_in_json():
    rc, d = last_json(["--steps", "1", "--warmup", "0"])
    assert "error" in d and "CUDA" in d["error"]


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_reference_arm_without_a_gpu_is_unavailable_not_a_crash():
    rc, d = last_json(["--impl", "reference", "--steps", "1", "--warmup", "0"])
    assert rc == 0 and d["impl"] == "reference"
    assert "unavailable" in d or d.get("unit") == "candidates/s"


def test_both_arms_quote_one_metric():
    sys.path.insert(0, str(ROOT))
    from baseline import contract

    src_ours = (ROOT / "bench.py").read_text()
    src_ref = (ROOT / "baseline" / "run_ref.py").read_text()
    assert "contract" in src_ours and "contract" in src_ref  # neither arm spells the metric out on its own
    assert contract.UNIT == "candidates/s" and "device time" in contract.METRIC
