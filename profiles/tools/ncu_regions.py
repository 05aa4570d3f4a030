#!/usr/bin/env python3
"""Where do a kernel's warps spend their time?  Reads the source page of an .ncu-rep (ncu --set full --import-source on) and prints
  * the stall-reason totals of the sampled warp states over the whole kernel, and
  * the same per barrier-delimited region of the SASS (a region ends at each BAR.SYNC / EXIT), with executed warp-instructions,
    sample share and the top stall reasons -- for the FFT kernels a region is one radix pass,
  * the hottest SASS lines.
usage: ncu_regions.py report.ncu-rep kernel-name-regex
"""
import collections
import csv
import io
import subprocess
import sys

rep, rx = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", "::regex:%s:1" % rx], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
print(rows[0][1][:140] if rows and len(rows[0]) > 1 else rows[0])
h = {k: i for i, k in enumerate(rows[1])}
stall_cols = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
seen, lines = set(), []
for r in rows[2:]:
    try:
        a, n, s = r[h["Address"]], int(r[h["Instructions Executed"]]), int(r[h["# Samples"]])
    except Exception:
        continue
    if a in seen:
        continue
    seen.add(a)
    lines.append((n, s, r[h["Source"]].strip(), {k: int(r[h[k]] or 0) for k in stall_cols}))
tot = sum(x[1] for x in lines)
allst = collections.Counter()
for _, _, _, st in lines:
    allst.update(st)
print("samples %d" % tot)
for k, v in allst.most_common(10):
    print("  %-26s %5d %5.1f%%" % (k, v, 100.0 * v / max(1, tot)))
reg, cur = [], [0, 0, collections.Counter(), 0]
for n, s, t, st in lines:
    cur[0] += n
    cur[1] += s
    cur[3] += 1
    cur[2].update(st)
    if "BAR.SYNC" in t or "EXIT" in t:
        reg.append(cur)
        cur = [0, 0, collections.Counter(), 0]
reg.append(cur)
for r in reg:
    if r[0] == 0 and r[1] == 0:
        continue
    top = ", ".join("%s %d" % (k.replace("stall_", ""), v) for k, v in r[2].most_common(4))
    print("region: %4d SASS lines  %9d warp-instr  %5d samples (%4.1f%%)  %s" % (r[3], r[0], r[1], 100.0 * r[1] / max(1, tot), top))
for n, s, t, st in sorted(lines, key=lambda x: -x[1])[:10]:
    print("%5d samples %8d exec  %s" % (s, n, t[:100]))
