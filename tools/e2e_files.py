#!/usr/bin/env python3
"""End-to-end timing of the drop-in CLI from FILES (host ingest / inflate included):
bin/select_STR_reads on a synthetic C2 prefix, plain and .gz, stdout to /dev/null, next to
the reference runner (oracle/ref_select.py -p <cores>) on a smaller prefix of the same data.
usage: tools/e2e_files.py [reads] [ref_reads]   -> one JSON line"""
import gzip, json, os, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from baitstr_b200 import synth

reads = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
ref_reads = int(sys.argv[2]) if len(sys.argv) > 2 else 200_000
tmp = tempfile.mkdtemp(prefix="bstr_e2e_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
cli = os.path.join(ROOT, "bin", "select_STR_reads")
res = {"reads": reads, "read_len": 100, "cores": os.cpu_count()}
try:
    half = reads // 2
    plain, gz = [], []
    for mate in (1, 2):
        p = os.path.join(tmp, "syn2_%d.fq" % mate)
        arr = synth.generate(2, 100, 0, half, mate)
        arr.tofile(p)
        plain.append(p)
        g = p + ".gz"
        subprocess.check_call("gzip -1 -c %s > %s" % (p, g), shell=True)
        gz.append(g)
    res["plain_bytes"] = sum(os.path.getsize(p) for p in plain)
    res["gz_bytes"] = sum(os.path.getsize(p) for p in gz)
    for tag, files in (("plain", plain), ("gz", gz)):
        best = None
        for _ in range(2):
            t0 = time.perf_counter()
            with open(os.devnull, "wb") as dn:
                subprocess.run([sys.executable, cli, "-p", str(os.cpu_count())] + files, stdout=dn, stderr=dn, check=True)
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        res["cli_%s_s" % tag] = round(best, 3)
        res["cli_%s_reads_per_s" % tag] = round(reads / best, 1)
        with open(os.devnull, "wb") as dn:       # one more run with the library's own breakdown (CUDA start-up excluded)
            tr = subprocess.run([sys.executable, cli, "-p", str(os.cpu_count())] + files, stdout=dn, stderr=subprocess.PIPE, check=True,
                                env=dict(os.environ, BSTR_TRACE="1")).stderr.decode()
        infl = [l[len("[bstr] "):] for l in tr.splitlines() if l.startswith("[bstr] inflate")]
        if infl:
            res["inflate_trace"] = infl
        line = [l for l in tr.splitlines() if l.startswith("[bstr] run_files")]
        if line:
            res["run_files_%s_trace" % tag] = line[0][len("[bstr] "):]
            ms = float(line[0].split()[2])
            res["run_files_%s_reads_per_s" % tag] = round(reads / (ms / 1e3), 1)
    # byte-compare against the reference on a prefix
    rh = ref_reads // 2
    small = []
    for mate in (1, 2):
        p = os.path.join(tmp, "small_%d.fq" % mate)
        synth.generate(2, 100, 0, rh, mate).tofile(p)
        small.append(p)
    t0 = time.perf_counter()
    ref = subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "ref_select.py"), "-p", str(os.cpu_count())] + small,
                         capture_output=True, check=True).stdout
    res["ref_s"] = round(time.perf_counter() - t0, 3)
    res["ref_reads_per_s"] = round(2 * rh / res["ref_s"], 1)
    ours = subprocess.run([sys.executable, cli] + small, capture_output=True, check=True).stdout
    res["byte_identical_to_reference"] = ours == ref
finally:
    for f in os.listdir(tmp):
        os.unlink(os.path.join(tmp, f))
    os.rmdir(tmp)
print(json.dumps(res))
