"""Times the GPU k-mer counting pass (bstr_kmers_*) against the reference's own
function (oracle/_ref/downstream/kmer_count_ref = extend_STR_reads.c:31-194 compiled in place) on reads
drawn from one random genome, checks that both give the same table, prints one JSON line.
  python tools/kmer_bench.py [nreads] [L] [genome_len] [k]"""
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from baitstr_b200 import capi                 # noqa: E402
from tests.util import genome_reads     # noqa: E402

nreads, L, glen, k = [int(a) for a in sys.argv[1:5]] + [1000000, 150, 3000000, 27][len(sys.argv) - 1:]
raw = genome_reads(nreads, L, glen, seed=5)
expected = nreads * (L - k + 1) // 2 + glen
with capi.Context() as ctx:
    ctx.count_kmers([raw[: len(raw) // 50]], k, expected_kmers=1 << 16 if nreads < 1000 else expected // 20)   # warm-up
    best = 1e9
    for _ in range(3):
        t = time.perf_counter()
        kmers, counts, st = ctx.count_kmers([raw], k, expected_kmers=expected)
        best = min(best, time.perf_counter() - t)
    # the same table from the batches the select stage leaves resident (bstr_kmers_add_resident): no second upload
    step = 64 << 20
    pinned = capi.PinnedBuffer(len(raw))
    import numpy as np
    pinned.array[:] = np.frombuffer(raw, np.uint8)
    from_pinned = 1e9                      # the same call with the reads in page-locked memory (bstr_host_alloc)
    for _ in range(3):
        t = time.perf_counter()
        ctx.count_kmers([pinned.array], k, expected_kmers=expected)
        from_pinned = min(from_pinned, time.perf_counter() - t)
    fused = 1e9
    for _ in range(3):
        t = time.perf_counter()
        with ctx.kmer_counter(k, expected) as kc:
            pos = 0
            while pos < len(raw):
                n = min(step, len(raw) - pos)
                out, info = ctx.scan_batch(pinned.array[pos:pos + n], final=pos + n >= len(raw))
                kc.add_resident(info["consumed"])
                pos += info["consumed"]
            k2, c2 = kc.finish()
        fused = min(fused, time.perf_counter() - t)
    select_only = 1e9
    for _ in range(3):
        t = time.perf_counter()
        pos = 0
        while pos < len(raw):
            n = min(step, len(raw) - pos)
            out, info = ctx.scan_batch(pinned.array[pos:pos + n], final=pos + n >= len(raw))
            pos += info["consumed"]
        select_only = min(select_only, time.perf_counter() - t)
    fused_same = bool((k2 == kmers).all() and (c2 == counts).all())
ref_bin = os.path.join(ROOT, "oracle", "_ref", "downstream", "kmer_count_ref")
ref_s, same = None, None
if os.path.exists(ref_bin):
    with tempfile.NamedTemporaryFile(suffix=".fq", dir="/dev/shm") as f:
        f.write(raw)
        f.flush()
        t = time.perf_counter()
        out = subprocess.run([ref_bin, str(k), "2", "255", str(expected), f.name], check=True, capture_output=True).stdout
        ref_s = time.perf_counter() - t
    want = [tuple(int(x) for x in line.split()) for line in out.decode().splitlines()]
    same = want == list(zip(kmers.tolist(), counts.tolist()))
print(json.dumps({"what": "k-mer counting pass of extend_STR_reads, host buffer in, sorted table out",
                  "reads": nreads, "read_len": L, "k": k, "kmers": st["kmers"], "distinct": st["distinct"],
                  "kept": int(len(kmers)), "gpu_s": round(best, 4), "gpu_reads_per_s": round(nreads / best),
                  "gpu_kmers_per_s": round(st["kmers"] / best),
                  "gpu_s_from_pinned_memory": round(from_pinned, 4),
                  "select_then_count_resident_s": round(fused, 4), "select_alone_s": round(select_only, 4), "resident_table_identical": fused_same,
                  "reference_s": None if ref_s is None else round(ref_s, 2),
                  "reference_reads_per_s": None if ref_s is None else round(nreads / ref_s),
                  "identical_to_reference": same}))
