#!/usr/bin/env python3
"""Phase attribution of k_fused_tile from an `ncu --page source --print-source cuda,sass --csv`
export: every SASS address is classified by the innermost source line of ours in its inline
stack.  usage: ncu_phases.py export.csv"""
import csv, sys
from collections import defaultdict
csv.field_size_limit(1 << 30)
RANGES = {
 'matcher.cuh': [(44,131,'match: bit helpers'),(143,245,'match: encode planes'),(262,373,'match: scan_period'),
                 (374,428,'match: pick winner'),(429,500,'match: match_read_bits'),(501,685,'match: long/dispatch'),(686,722,'validate bases')],
 'kernels_v1.cuh': [(32,37,'index: nl_flags'),(210,228,'outlen'),(288,304,'format: digits/upper/comp')],
 'fused.cuh': [(85,107,'tma/mbarrier'),(109,149,'lookback'),(151,159,'index: nl_mask'),(161,183,'block scan'),(194,203,'record lines'),(204,277,'format: copy'),
               (279,312,'B index chunks'),(319,360,'A stage'),(361,387,'B index fill'),(388,453,'C record phase + bookkeeping'),(454,502,'D per-record glue'),
               (503,520,'E offsets'),(521,551,'F format glue + lookback2'),(552,580,'G flush')],
}
PRIO = ['matcher.cuh', 'kernels_v1.cuh', 'fused.cuh']
fname = cols = cur = None; seen = []; stack = defaultdict(dict); inst = {}; samp = {}; thr = {}; stalls = defaultdict(lambda: defaultdict(float))
for row in csv.reader(open(sys.argv[1])):
    if not row: continue
    if row[0] == "File Path":
        fname = row[1].split('/')[-1]
        if seen and fname == seen[0]: break
        seen.append(fname); cols = None; continue
    if row[0] == "Line No":
        cols = row; ci = cols.index("Instructions Executed"); cs = cols.index("# Samples"); ct = cols.index("Thread Instructions Executed")
        sc = [(i, c) for i, c in enumerate(cols) if c.startswith("stall_") and "Not Issued" not in c]; continue
    if cols is None or row[0] == "Function Name": continue
    if row[0].strip().isdigit(): cur = int(row[0])
    if len(row) > 2 and row[2].strip().startswith('0x'):
        a = row[2].strip(); stack[a].setdefault(fname, cur)
        try:
            inst[a] = float(row[ci] or 0); samp[a] = float(row[cs] or 0); thr[a] = float(row[ct] or 0)
            for i, c in sc:
                if row[i]: stalls[a][c] = float(row[i])
        except ValueError: pass
agg = defaultdict(lambda: [0, 0, 0, defaultdict(float)])
for a, v in inst.items():
    name = 'other'
    for f in PRIO:
        l = stack[a].get(f)
        if l is None: continue
        hit = [nm for lo, hi, nm in RANGES[f] if lo <= l <= hi]
        if hit: name = hit[0]; break
    g = agg[name]; g[0] += v; g[1] += samp[a]; g[2] += thr[a]
    for c, x in stalls[a].items(): g[3][c] += x
ti = sum(inst.values()); ts = sum(samp.values())
print("SASS instructions %d, warp-inst executed %.0f, samples %.0f" % (len(inst), ti, ts))
for nm, (i, s, t, st) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    top = sorted(st.items(), key=lambda kv: -kv[1])[:3]
    print("%-28s inst %5.1f%%  samples %5.1f%%  lanes/inst %4.1f   %s" % (nm, 100*i/ti, 100*s/ts, t/max(i,1), " ".join("%s=%.0f" % (k.replace('stall_',''), x) for k, x in top)))
