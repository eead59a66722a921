#!/usr/bin/env python3
"""Phase attribution of k_fused_tile from an `ncu --page source --print-source cuda,sass --csv`
export: every SASS address is classified by the innermost source line of ours in its inline
stack; lines are mapped to the function that contains them (and, inside k_fused_tile, to the
phase comment `// ---- X: ...` above them), read from the CURRENT sources.
usage: ncu_phases.py export.csv [--lines]"""
import csv, os, re, sys
from collections import defaultdict
csv.field_size_limit(1 << 30)
SRC = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baitstr_b200", "csrc")
FUNC = re.compile(r"^(?:template\s*<[^>]*>\s*)?(?:__device__|__global__|static|inline|__host__|__forceinline__|\s)+[\w:<>\*&\s]*?\b(\w+)\s*\([^;]*$")
GROUP = [  # function name (regex) -> report group
    (r"first_trip|extend_trips|scan_period", "match: period scan"), (r"WarpWork|push|run|finish|reset_keys|match_read_warp|match_read_tile", "match: work list"),
    (r"pick_|motif_is|scan_finish|take_if", "match: pick winner"), (r"encode_window|gather_top|mark_invalid", "match: encode planes"),
    (r"scan_first|match_read", "match: scan_first glue"), (r"shr|shl|and_shr|lowmask|any|extract32|highest|lowest|testbit", "match: bit helpers"),
    (r"swar_|bases_invalid", "validate bases"), (r"nl_flags|nl_mask", "index: nl flags/mask"), (r"index_chunks", "B index chunks"),
    (r"block_excl_scan", "block scan"), (r"lookback", "lookback"), (r"mbar|bulk_load|smem_addr", "tma/mbarrier"),
    (r"copy_xform|xform_", "format: copy"), (r"format_record|put_dec|upper|comp_base", "format: digits/glue"), (r"dec_digits|record_out_len", "outlen"),
    (r"tile_record_lines", "record lines"),
]
def build():
    rng = {}
    for f in ("matcher.cuh", "kernels_v1.cuh", "fused.cuh"):
        lines = open(os.path.join(SRC, f), errors="replace").read().split("\n")
        labels, cur, phase, depth, stmt, fn_depth = [], "other", None, 0, "", None
        for no, ln in enumerate(lines, 1):
            code = ln.split("//")[0]
            if fn_depth is None:
                stmt += " " + code.strip()
                if "{" in code:
                    head = stmt.split("{")[0]
                    head = re.sub(r"__launch_bounds__\([^;{]*?\)\s*(?=k_)", "", head)
                    m = re.search(r"(\w+)\s*\((?:[^()]|\([^()]*\))*\)\s*(?:const)?\s*$", head.strip())
                    if m and not re.match(r"\s*(namespace|struct|enum|class|extern)\b", stmt.strip()) and m.group(1) not in ("if", "for", "while", "switch"):
                        cur, fn_depth, phase = m.group(1), depth, None
                    stmt = ""
                elif code.strip().endswith((";", "}")) or not code.strip():
                    stmt = "" if code.strip().endswith((";", "}")) else stmt
            if cur.startswith("k_fused_tile") and fn_depth is not None:
                m = re.match(r"\s*// ---- ([A-G]):? ?(.*)", ln)
                if m: phase = "%s %s" % (m.group(1), " ".join(m.group(2).split()[:4]))
            labels.append((cur if fn_depth is not None else "other", phase if (fn_depth is not None and cur.startswith("k_fused_tile")) else None))
            depth += code.count("{") - code.count("}")
            if fn_depth is not None and depth <= fn_depth:
                fn_depth = None; cur = "other"; stmt = ""
        rng[f] = labels
    return rng
def group_of(fn, phase):
    if phase: return "kernel " + phase
    for pat, g in GROUP:
        if re.fullmatch(pat, fn) or re.match("(?:" + pat + ")$", fn): return g
    return "fn " + fn
RNG = build()
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
byline = defaultdict(lambda: [0, 0, 0])
for a, v in inst.items():
    name = 'other'
    for f in PRIO:
        l = stack[a].get(f)
        if l is None or l > len(RNG[f]): continue
        fn, ph = RNG[f][l - 1]
        name = group_of(fn, ph)
        b = byline[(f, l)]; b[0] += v; b[1] += samp[a]; b[2] += thr[a]
        break
    g = agg[name]; g[0] += v; g[1] += samp[a]; g[2] += thr[a]
    for c, x in stalls[a].items(): g[3][c] += x
ti = sum(inst.values()); ts = sum(samp.values())
print("SASS instructions %d, warp-inst executed %.0f, samples %.0f" % (len(inst), ti, ts))
for nm, (i, s, t, st) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    top = sorted(st.items(), key=lambda kv: -kv[1])[:3]
    print("%-34s inst %5.1f%%  samples %5.1f%%  lanes/inst %4.1f   %s" % (nm, 100*i/ti, 100*s/ts, t/max(i,1), " ".join("%s=%.0f" % (k.replace('stall_',''), x) for k, x in top)))
if "--lines" in sys.argv:
    print("-- hottest source lines (samples)")
    for (f, l), (i, s, t) in sorted(byline.items(), key=lambda kv: -kv[1][1])[:40]:
        src = open(os.path.join(SRC, f), errors="replace").read().split("\n")[l - 1].strip()[:90]
        print("%-14s %4d  inst %4.1f%% samples %4.1f%% lanes %4.1f  %s" % (f, l, 100*i/ti, 100*s/ts, t/max(i,1), src))
