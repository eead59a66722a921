"""Times the GPU merge stage (bstr_merge_reads) against the reference's own
merge_STR_reads (oracle/_ref/downstream, its source compiled in place) on synthetic STR loci: reads ->
this repository's select stage on the GPU -> merge.  Checks that both print the same bytes; one JSON line.
  BSTR_TRACE=1 python tools/merge_bench.py [nloci] [coverage] [read_len]"""
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from baitstr_b200 import capi     # noqa: E402
from tests import util            # noqa: E402

nloci, cov, L = [int(a) for a in sys.argv[1:4]] + [4000, 40, 150][len(sys.argv) - 1:]
raw = b"".join(util.str_locus_reads(500, cov, L, seed=100 + i) .replace(b"@m.", b"@m%d." % i) for i in range((nloci + 499) // 500))
with capi.Context() as ctx:
    sel = b""
    step = 64 << 20
    pos = 0
    while pos < len(raw):
        out, info = ctx.scan_batch(raw[pos:pos + step], final=pos + step >= len(raw))
        sel += out
        pos += info["consumed"]
    nsel = sel.count(b"\n") // 4
    ctx.merge_reads(sel[: sel.index(b"\n@", len(sel) // 50) + 1], 27, 4, 10000, False)       # warm-up
    best = 1e9
    for _ in range(3):
        t = time.perf_counter()
        got, st = ctx.merge_reads(sel, 27, 4, 10000, False)
        best = min(best, time.perf_counter() - t)
ref_bin = os.path.join(ROOT, "oracle", "_ref", "downstream", "merge_STR_reads")
ref_s, same = None, None
if os.path.exists(ref_bin):
    with tempfile.NamedTemporaryFile(suffix=".fq", dir="/dev/shm") as f:
        f.write(sel)
        f.flush()
        t = time.perf_counter()
        want = subprocess.run([ref_bin, "27", f.name], check=True, capture_output=True).stdout
        ref_s = time.perf_counter() - t
    same = want == got
print(json.dumps({"what": "merge stage: select output in host memory -> printed blocks", "loci": nloci, "coverage": cov, "read_len": L,
                  "reads_in": raw.count(b"\n") // 4, "reads_selected": nsel, **st, "gpu_s": round(best, 4),
                  "gpu_selected_reads_per_s": round(nsel / best), "reference_s": None if ref_s is None else round(ref_s, 2),
                  "reference_selected_reads_per_s": None if ref_s is None else round(nsel / ref_s), "identical_to_reference": same}))
