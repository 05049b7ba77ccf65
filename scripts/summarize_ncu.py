#!/usr/bin/env python
"""Turn an .ncu-rep (brought back in gpurun_out/) into a small text summary under profiles/.

    python scripts/summarize_ncu.py gpurun_out/prof_pair.ncu-rep profiles/ncu_pair_x256.txt "note"
"""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "sm__cycles_elapsed.avg.per_second", "launch__grid_size", "launch__block_size",
    "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.per_cycle_active", "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
]


def main():
    rep, out, note = sys.argv[1], sys.argv[2], (sys.argv[3] if len(sys.argv) > 3 else "")
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    lines = [f"# {rep}", f"# {note}", "# ncu --set full --clock-control none --import-source on (one kernel launch per row)", ""]
    for vals in rows[2:]:
        name = vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
        lines.append(f"## kernel: {name}")
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                lines.append(f"{k:95s} {vals[i]:>18s} {units[i]}")
        lines.append("")
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    srows = list(csv.reader(io.StringIO(src)))
    if len(srows) > 2:
        h = srows[1]
        ia, isrc, iex, ismp = h.index("Address"), h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
        data = [r for r in srows[2:] if len(r) > iex and r[ia].startswith("0x")]
        tot_s = sum(int(r[ismp] or 0) for r in data) or 1
        tot_e = sum(int(r[iex] or 0) for r in data) or 1
        ops = {}
        for r in data:
            w = r[isrc].split()
            if not w:
                continue
            op = w[1] if w[0].startswith("@") and len(w) > 1 else w[0]
            e = ops.setdefault(op, [0, 0])
            e[0] += int(r[iex] or 0)
            e[1] += int(r[ismp] or 0)
        lines.append("## executed warp-instructions by SASS opcode (top 16) — share of instructions / share of stall samples")
        for op, (e, s) in sorted(ops.items(), key=lambda kv: -kv[1][0])[:16]:
            lines.append(f"{op:28s} {e:14d} {100 * e / tot_e:6.2f}%   samples {100 * s / tot_s:6.2f}%")
        lines.append("")
        lines.append("## 12 hottest instructions by stall samples")
        for r in sorted(data, key=lambda r: -int(r[ismp] or 0))[:12]:
            lines.append(f"{r[ia]}  {100 * int(r[ismp] or 0) / tot_s:5.2f}%  {r[isrc][:90]}")
    open(out, "w").write("\n".join(lines) + "\n")
    print("wrote", out)


if __name__ == "__main__":
    main()
