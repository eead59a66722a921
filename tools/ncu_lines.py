#!/usr/bin/env python3
"""Aggregate an `ncu --page source --print-source cuda,sass --csv` export per CUDA source line.
usage: ncu_lines.py export.csv [topN] [kernel-instance-index]"""
import csv
import sys
from collections import defaultdict

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
csv.field_size_limit(1 << 30)
inst = defaultdict(float); samp = defaultdict(float); text = {}
stall = defaultdict(lambda: defaultdict(float))
fname = None; cols = None; kernels = 0; cur_line = None
for row in csv.reader(open(path)):
    if not row:
        continue
    if row[0] == "Kernel Name":
        kernels += 1
        continue
    if kernels > 1:
        break
    if row[0] == "File Name":
        fname = row[1].split("/")[-1]; cols = None
        continue
    if row[0] == "Line No":
        cols = row
        ci = cols.index("Instructions Executed"); cs = cols.index("# Samples")
        stall_cols = [(i, c) for i, c in enumerate(cols) if c.startswith("stall_") and "Not Issued" not in c]
        continue
    if cols is None:
        continue
    if row[0].strip() and not row[0].strip().isdigit():
        continue
    if row[0].strip():
        cur_line = (fname, int(row[0])); text[cur_line] = row[1].strip()[:90]
    if len(row) > ci and row[2].strip():
        try:
            inst[cur_line] += float(row[ci] or 0); samp[cur_line] += float(row[cs] or 0)
            for i, c in stall_cols:
                if row[i]:
                    stall[cur_line][c] += float(row[i])
        except ValueError:
            pass
tot = sum(inst.values()); ts = sum(samp.values())
print("total warp-instructions %.0f, samples %.0f" % (tot, ts))
byfile = defaultdict(float)
for k, v in inst.items():
    byfile[k[0]] += v
print({k: "%.1f%%" % (100 * v / tot) for k, v in byfile.items()})
for k, v in sorted(inst.items(), key=lambda kv: -kv[1])[:top]:
    st = sorted(stall[k].items(), key=lambda kv: -kv[1])[:2]
    print("%5.1f%% inst %5.1f%% samp  %s:%d  %s   %s" % (100 * v / tot, 100 * samp[k] / max(ts, 1), k[0], k[1], text.get(k, ""),
          " ".join("%s=%.0f" % (a.replace("stall_", ""), b) for a, b in st)))
