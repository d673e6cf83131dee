#!/usr/bin/env python
"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump per CUDA source line.
usage: ncu -i rep.ncu-rep --page source --csv --print-source cuda,sass > x.csv; python scripts/ncu_lines.py x.csv [top]
The dump is a sequence of sections (one per kernel x source file): `File Path`, `Function Name`,
a `Line No` header, then one summary row per source line followed by its SASS rows."""
import csv
import sys
from collections import defaultdict

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
kernel, fname, hdr = None, None, None
agg = defaultdict(dict)                      # kernel -> (file, line) -> [inst, samples, text]
for r in rows:
    if not r:
        continue
    if r[0] in ("Kernel Name", "Function Name"):
        kernel = r[1]; continue
    if r[0] in ("File Name", "File Path"):
        fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No":
        hdr = r; ii = hdr.index("Instructions Executed"); si = hdr.index("# Samples"); continue
    if hdr is None or r[0] == "" or len(r) <= ii or not r[ii].isdigit():
        continue
    agg[kernel][(fname, r[0])] = [int(r[ii]), int(r[si]) if r[si].isdigit() else 0, r[1].strip()[:90]]
for k, d in agg.items():
    tot = sum(v[0] for v in d.values()); ts = sum(v[1] for v in d.values()) or 1
    if not tot:
        continue
    print("==", str(k)[:100], "warp-inst", tot)
    for (f, ln), v in sorted(d.items(), key=lambda kv: -kv[1][0])[:top]:
        print("  %5.1f%% inst %5.1f%% smp  %s:%s  %s" % (100 * v[0] / tot, 100 * v[1] / ts, f, ln, v[2]))
