#!/usr/bin/env python3
"""Aggregate an `ncu -i X.ncu-rep --page source --print-source cuda,sass --csv` export per CUDA
source line (first kernel instance only).

ncu attributes an inlined SASS instruction to every level of its inline stack, so a line of
the kernel body carries the cost of everything it calls; `--body FILE:LO-HI` restricts the
table to the kernel body and therefore gives a per-phase attribution that sums to 100 %.

usage: ncu_lines.py export.csv [topN] [--file NAME] [--body NAME:LO-HI]
"""
import csv
import sys
from collections import defaultdict

args = [a for a in sys.argv[1:] if not a.startswith("--")]
path = args[0]
top = int(args[1]) if len(args) > 1 else 40
only_file = body = None
for i, a in enumerate(sys.argv):
    if a == "--file":
        only_file = sys.argv[i + 1]
    if a == "--body":
        f, r = sys.argv[i + 1].split(":")
        lo, hi = r.split("-")
        body = (f, int(lo), int(hi))
args = [a for a in args if a not in (only_file, body and sys.argv[sys.argv.index("--body") + 1])]

csv.field_size_limit(1 << 30)
inst = defaultdict(float); samp = defaultdict(float); text = {}
stall = defaultdict(lambda: defaultdict(float))
fname = None; cols = None; cur_line = None; seen = set()
for row in csv.reader(open(path)):
    if not row:
        continue
    if row[0] in ("File Path", "File Name"):
        fname = row[1].split("/")[-1]; cols = None
        if fname in seen and fname == first_file:
            break                                   # second kernel instance starts
        if not seen:
            first_file = fname
        seen.add(fname)
        continue
    if row[0] == "Line No":
        cols = row
        ci = cols.index("Instructions Executed"); cs = cols.index("# Samples")
        stall_cols = [(i, c) for i, c in enumerate(cols) if c.startswith("stall_") and "Not Issued" not in c]
        continue
    if cols is None or row[0] in ("Function Name", "Kernel Name"):
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


def keep(k):
    if body:
        return k[0] == body[0] and body[1] <= k[1] <= body[2]
    if only_file:
        return k[0] == only_file
    return True


keys = [k for k in inst if keep(k)]
tot = sum(inst[k] for k in keys); ts = sum(samp[k] for k in keys)
print("warp-instructions %.0f, samples %.0f (%s)" % (tot, ts, "body " + str(body) if body else only_file or "all files, inline stacks counted at every level"))
for k in sorted(keys, key=lambda k: -samp[k] if body else -inst[k])[:top]:
    st = sorted(stall[k].items(), key=lambda kv: -kv[1])[:3]
    print("%5.1f%% inst %5.1f%% samp  %s:%d  %s   %s" % (100 * inst[k] / tot, 100 * samp[k] / max(ts, 1), k[0], k[1], text.get(k, ""),
          " ".join("%s=%.0f" % (a.replace("stall_", ""), b) for a, b in st)))
