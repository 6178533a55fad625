#!/usr/bin/env python
"""Per-source-line instruction and stall-sample shares of one kernel from an ncu report captured with
`--set full --import-source on` (binary built with -lineinfo).
Usage: python tools/ncu_lines.py <file.ncu-rep> [top] [samples]   (third argument: sort by stall samples)"""
import csv, subprocess, sys, collections
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Line No"][0]
hdr = rows[hi]
ci, si = hdr.index("Instructions Executed"), hdr.index("# Samples")
data = []
for r in rows[hi + 1:]:
    if r and r[0].isdigit():
        try:
            data.append((int(r[ci]), int(r[si]), int(r[0]), r[1][:110]))
        except ValueError:
            pass
ti, ts = sum(d[0] for d in data), max(1, sum(d[1] for d in data))
print("warp instructions %d, stall samples %d" % (ti, ts))
by_samples = len(sys.argv) > 3
for ie, sm, ln, src in sorted(data, key=lambda d: (d[1], d[0]) if by_samples else d, reverse=True)[:top]:
    print("%5.1f%% instr %5.1f%% samples  L%-4d %s" % (100 * ie / ti, 100 * sm / ts, ln, src))
