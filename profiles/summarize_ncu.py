#!/usr/bin/env python3
"""Turn an .ncu-rep (ncu --set full --import-source on) into the small text summary kept under profiles/.

usage: summarize_ncu.py report.ncu-rep [kernel-regex] > profiles/rNN_<what>.txt
Prints, per captured launch, the metrics DESIGN.md quotes (duration, registers, occupancy limits, issue
utilisation, stall reasons per issue, DRAM / L2 / shared-memory traffic) and, for the first launch matching
the regex, the executed-instruction opcode histogram and the hottest sampled SASS lines.
"""
import collections
import csv
import io
import re
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum",
    "launch__grid_size",
    "launch__block_size",
    "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic",
    "launch__occupancy_limit_shared_mem",
    "launch__occupancy_limit_registers",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "lts__t_bytes.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
]


def ncu_csv(rep, page, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv", *extra], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    rx = sys.argv[2] if len(sys.argv) > 2 else "."
    rows = ncu_csv(rep, "raw")
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    print(f"# {rep}: per-launch metrics (ncu --set full --clock-control none)")
    for r in rows[2:]:
        name = r[ix["Kernel Name"]]
        short = re.sub(r"\(.*", "", name).replace("<unnamed>::", "")
        print(f"\n== {short}")
        for m in METRICS:
            if m in ix:
                print(f"  {m:86s} {r[ix[m]]:>16s} {units[ix[m]]}")
    src = ncu_csv(rep, "source", ["--kernel-id", f"::regex:{rx}:1"])
    if len(src) < 3:
        return
    print(f"\n# executed warp-instruction histogram, first launch matching /{rx}/: {src[0][1][:110]}")
    h = {k: i for i, k in enumerate(src[1])}
    ops, samp, lines, seen = collections.Counter(), collections.Counter(), [], set()
    for r in src[2:]:
        try:
            addr, n, s = r[h["Address"]], int(r[h["Instructions Executed"]]), int(r[h["# Samples"]])
        except (ValueError, KeyError, IndexError):
            continue
        if addr in seen:
            continue
        seen.add(addr)
        text = r[h["Source"]].strip()
        op = re.sub(r"^@!?U?P\d+\s+", "", text).split()[0].split(".")[0]
        ops[op] += n
        samp[op] += s
        lines.append((s, n, text))
    tot, stot = sum(ops.values()) or 1, sum(samp.values()) or 1
    print(f"total warp instructions {tot}, stall samples {stot}")
    for o, n in ops.most_common(24):
        print(f"  {o:10s} {n:12d} {100 * n / tot:5.1f}%   samples {100 * samp[o] / stot:5.1f}%")
    print("# hottest sampled SASS lines (samples, executions, text)")
    for s, n, text in sorted(lines, reverse=True)[:16]:
        print(f"  {s:6d} {n:10d}  {text[:110]}")


if __name__ == "__main__":
    main()
