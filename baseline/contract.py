"""The strings both arms of bench.py must agree on (the driver compares them verbatim)."""
METRIC = "candidates/s (one full solve of the input; device time of the solve's kernels, max over ranks)"
UNIT = "candidates/s"
MODEL = "exhaustive (n,k) alignment search"
