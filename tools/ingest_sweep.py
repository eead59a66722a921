#!/usr/bin/env python3
"""Sweep the host-side ingest knobs of bstr_run_files on one synthetic C2 prefix (plain and .gz):
usage: tools/ingest_sweep.py [reads]  -> one line per configuration (wall seconds of the CLI,
run_files milliseconds from BSTR_TRACE).  Knobs are the BSTR_* environment variables that
stream_host.cu reads (batch size, reader / decoder / filler threads, decode-ahead)."""
import os, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from baitstr_b200 import synth

reads = int(sys.argv[1]) if len(sys.argv) > 1 else 20_000_000
tmp = tempfile.mkdtemp(prefix="bstr_sweep_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
cli = os.path.join(ROOT, "bin", "select_STR_reads")
cores = os.cpu_count()
plain, gz = [], []
try:
    for mate in (1, 2):
        p = os.path.join(tmp, "syn2_%d.fq" % mate)
        synth.generate(2, 100, 0, reads // 2, mate).tofile(p)
        plain.append(p)
        subprocess.check_call("gzip -1 -c %s > %s.gz" % (p, p), shell=True)
        gz.append(p + ".gz")
    MB = 1 << 20
    configs = [("plain", {}), ("plain", {"BSTR_BATCH_BYTES": 16 * MB}), ("plain", {"BSTR_BATCH_BYTES": 4 * MB}),
               ("gz", {}), ("gz", {"BSTR_GZ_DECODERS": 13, "BSTR_GZ_FILLERS": 3}), ("gz", {"BSTR_GZ_DECODERS": 14, "BSTR_GZ_FILLERS": 2}),
               ("gz", {"BSTR_GZ_DECODERS": 14, "BSTR_GZ_FILLERS": 3}), ("gz", {"BSTR_BATCH_BYTES": 16 * MB}),
               ("gz", {"BSTR_GZ_PENDING": 3})]
    for kind, env in configs:
        files = plain if kind == "plain" else gz
        e = dict(os.environ, BSTR_TRACE="1", **{k: str(v) for k, v in env.items()})
        best, trace, infl = None, "", ""
        for _ in range(2):
            t0 = time.perf_counter()
            with open(os.devnull, "wb") as dn:
                r = subprocess.run([sys.executable, cli, "-p", str(cores)] + files, stdout=dn, stderr=subprocess.PIPE, env=e, check=True)
            dt = time.perf_counter() - t0
            if best is None or dt < best:
                best = dt
                trace = [l for l in r.stderr.decode().splitlines() if l.startswith("[bstr] run_files")][0]
                infl = " ".join(l[l.index("| threads:"):] for l in r.stderr.decode().splitlines() if l.startswith("[bstr] inflate"))
        print("%-5s %-90s wall %.2f s | %s" % (kind, env, best, trace[7:trace.index("| writer")] + (infl if kind == "gz" else "")), flush=True)
finally:
    for f in os.listdir(tmp):
        os.unlink(os.path.join(tmp, f))
    os.rmdir(tmp)
