"""The whole BaitSTR pipeline of the reference's README (select -> merge -> extend) through this repository's
three command lines (bin/, all work on the GPU) beside the reference's own three programs (oracle/_ref: its select
stage under Python 3 with -p <all cores>; oracle/_ref/downstream: its merge_STR_reads and extend_STR_reads
compiled in place) on reads of a synthetic diploid genome.  cmp of every stage's output; one JSON line.
  python tools/pipeline_bench.py [genome_len] [nloci] [coverage] [read_len]"""
import json
import os
import shutil
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import util            # noqa: E402

glen, nloci, cov, L = [int(a) for a in sys.argv[1:5]] + [10000000, 10000, 30, 150][len(sys.argv) - 1:]
cores = len(os.sched_getaffinity(0))
tmp = tempfile.mkdtemp(prefix="bstr_pipe_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
try:
    raw = util.diploid_str_reads_np(glen, nloci, cov, L, seed=41)
    nreads = raw.count(b"\n") // 4
    recs = raw.split(b"\n@")
    reads = [os.path.join(tmp, "reads_%d.fq" % i) for i in (1, 2)]
    with open(reads[0], "wb") as f:
        f.write(b"\n@".join(recs[:len(recs) // 2]) + b"\n")
    with open(reads[1], "wb") as f:
        f.write(b"@" + b"\n@".join(recs[len(recs) // 2:]))

    def run(cmd, out):
        t = time.perf_counter()
        with open(out, "wb") as f, open(os.devnull, "wb") as dn:
            subprocess.run(cmd, stdout=f, stderr=dn, check=True)
        return time.perf_counter() - t

    py, b = sys.executable, lambda n: os.path.join(ROOT, "bin", n)
    down = os.path.join(ROOT, "oracle", "_ref", "downstream")
    o = lambda n: os.path.join(tmp, n)
    run([py, b("select_STR_reads"), "-n", "3"] + reads, o("warm.fq"))                    # first CUDA use of the box
    ours = {"select": run([py, b("select_STR_reads"), "-n", "3"] + reads, o("sel.fq")),
            "merge": run([py, b("merge_STR_reads"), "27", o("sel.fq")], o("merged.fq")),
            "extend": run([py, b("extend_STR_reads"), str(glen), str(cov), "27", o("merged.fq")] + reads, o("contigs.fa"))}
    ref, same = None, None
    if os.path.exists(os.path.join(down, "merge_STR_reads")) and os.path.isdir(os.path.join(ROOT, "oracle", "_ref")):
        ref = {"select": run([py, os.path.join(ROOT, "oracle", "ref_select.py"), "-p", str(cores), "-n", "3"] + reads, o("r_sel.fq")),
               "merge": run([os.path.join(down, "merge_STR_reads"), "27", o("r_sel.fq")], o("r_merged.fq")),
               "extend": run([os.path.join(down, "extend_STR_reads"), str(glen), str(cov), "27", o("r_merged.fq")] + reads, o("r_contigs.fa"))}
        same = {k: open(o(a), "rb").read() == open(o(c), "rb").read() for k, a, c in
                (("select", "sel.fq", "r_sel.fq"), ("merge", "merged.fq", "r_merged.fq"), ("extend", "contigs.fa", "r_contigs.fa"))}
    count = lambda p, c: open(o(p), "rb").read().count(c)
    print(json.dumps({"what": "select_STR_reads -n 3 | merge_STR_reads 27 | extend_STR_reads gs cov 27: command lines, files in tmpfs, process start-up included",
                      "genome": glen, "loci": nloci, "reads": nreads, "read_len": L, "selected": count("sel.fq", b"\n") // 4,
                      "blocks": count("merged.fq", b"@Block"), "contigs": count("contigs.fa", b">"),
                      "seconds": {k: round(v, 3) for k, v in ours.items()}, "total_s": round(sum(ours.values()), 3),
                      "reference_seconds": None if ref is None else {k: round(v, 2) for k, v in ref.items()},
                      "reference_total_s": None if ref is None else round(sum(ref.values()), 2), "reference_cores_select": cores,
                      "identical_to_reference": same}))
finally:
    shutil.rmtree(tmp, ignore_errors=True)
