"""Times the GPU extend stage (bstr_kmers_* + bstr_kmers_extend) against the reference's own
extend_STR_reads (oracle/_ref/downstream, its source compiled in place) on a synthetic diploid genome:
reads -> this repository's select and merge stages on the GPU -> extend.  Checks that both print the same
bytes; one JSON line.
  BSTR_TRACE=1 python tools/extend_bench.py [genome_len] [nloci] [coverage] [read_len]"""
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

glen, nloci, cov, L = [int(a) for a in sys.argv[1:5]] + [500000, 500, 30, 150][len(sys.argv) - 1:]
raw = util.diploid_str_reads(glen, nloci, cov, L, seed=21)
nreads = raw.count(b"\n") // 4
recs = raw.split(b"\n@")
r1, r2 = b"\n@".join(recs[:len(recs) // 2]) + b"\n", b"@" + b"\n@".join(recs[len(recs) // 2:])
with capi.Context(n=3) as ctx:
    sel = ctx.scan_batch(raw)[0]
    merged, mst = ctx.merge_reads(sel, 27, 3, 10000, True)
    expected = nreads * (L - 26) // 2 + 2 * glen
    ctx.extend_reads(merged, [r1[: r1.index(b"\n@", len(r1) // 20) + 1]], 27, expected_kmers=expected // 10)      # warm-up
    best = 1e9
    for _ in range(3):
        t = time.perf_counter()
        got, st = ctx.extend_reads(merged, [r1, r2], 27, expected_kmers=expected)
        best = min(best, time.perf_counter() - t)
ref_bin = os.path.join(ROOT, "oracle", "_ref", "downstream", "extend_STR_reads")
ref_s, same = None, None
if os.path.exists(ref_bin):
    d = tempfile.mkdtemp(dir="/dev/shm")
    paths = []
    for fn, data in (("merged.fq", merged), ("r1.fq", r1), ("r2.fq", r2)):
        with open(os.path.join(d, fn), "wb") as f:
            f.write(data)
        paths.append(os.path.join(d, fn))
    t = time.perf_counter()
    want = subprocess.run([ref_bin, str(glen), str(cov), "27"] + paths, check=True, capture_output=True).stdout
    ref_s = time.perf_counter() - t
    same = want == got
    for p in paths:
        os.unlink(p)
    os.rmdir(d)
print(json.dumps({"what": "extend stage: merged blocks + reads in host memory -> contigs (k-mer counting pass + walks)", "genome": glen, "loci": nloci,
                  "reads": nreads, "read_len": L, "selected": sel.count(b"\n") // 4, **st, "longest_contig": max(len(x) for x in got.split(b"\n")),
                  "gpu_s": round(best, 4), "gpu_reads_per_s": round(nreads / best), "reference_s": None if ref_s is None else round(ref_s, 2),
                  "reference_reads_per_s": None if ref_s is None else round(nreads / ref_s), "identical_to_reference": same}))
