#!/usr/bin/env python3
"""profiles/r02_traffic.json from the `ncu --set full` captures of bench.py (run here, on the .ncu-rep files gpurun brought back):
DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) of the acquisition inverse pair and of the tracking kernel,
which bench.py reports as roofline.traffic.

usage: traffic_from_ncu.py <acq_inverse.ncu-rep> <tracking.ncu-rep> > profiles/r02_traffic.json
"""
import csv
import io
import json
import subprocess
import sys


def rows(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(io.StringIO(out)))
    hdr = r[0]
    units = r[1]
    ix = {h: i for i, h in enumerate(hdr)}

    def val(row, name):
        v = float(row[ix[name]].replace(",", ""))
        u = units[ix[name]].lower()
        return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)

    return [(row[ix["Kernel Name"]], val(row, "dram__bytes_read.sum") + val(row, "dram__bytes_write.sum"), val(row, "gpu__time_duration.sum"),
             row[ix["launch__grid_size"]]) for row in r[2:]]


def main():
    acq, trk = rows(sys.argv[1]), rows(sys.argv[2])
    row = [a for a in acq if "row" in a[0] or "fused" in a[0]]
    col = [a for a in acq if "col" in a[0]]
    pair = (sum(a[1] for a in row) / max(1, len(row))) + (sum(a[1] for a in col) / max(1, len(col)))
    t = [a for a in trk if "gb_trk_kernel" in a[0]]
    out = {"acq_pair_dram_bytes": pair, "trk_dram_bytes": sum(a[1] for a in t) / max(1, len(t)),
           "source": "profiles/tools/traffic_from_ncu.py on %s (%d row + %d column launches, grids %s / %s) and %s (%d launches)" % (
               sys.argv[1].split("/")[-1], len(row), len(col), row[0][3] if row else "-", col[0][3] if col else "-", sys.argv[2].split("/")[-1], len(t)),
           "launches": {"acq": [{"kernel": a[0][:60], "dram_bytes": a[1], "ns": a[2], "grid": a[3]} for a in acq],
                        "trk": [{"kernel": a[0][:60], "dram_bytes": a[1], "ns": a[2], "grid": a[3]} for a in t]}}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
