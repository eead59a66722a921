#!/usr/bin/env python3
"""Stage timings of the GPU inflater against the host inflater on a gzip -1 file of synthetic C2 reads."""
import os, subprocess, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from baitstr_b200 import capi, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
tmp = tempfile.mkdtemp(prefix="gib_", dir="/dev/shm")
p = os.path.join(tmp, "r.fq")
synth.generate(2, 100, 0, n, 1).tofile(p)
subprocess.check_call("gzip -1 -k %s" % p, shell=True)
os.environ["BSTR_TRACE"] = "1"
for win, rng in ((0, 0), (256 << 20, 64 << 10), (256 << 20, 256 << 10), (64 << 20, 64 << 10)):
    for rep in range(2):
        t0 = time.perf_counter(); st = capi.gunzip_gpu(p + ".gz", -1, 0, win, rng); dt = time.perf_counter() - t0
    print("gpu  window %3d MiB range %3d KiB: %.1f ms  %.2f GB/s inflated  %s" % (win >> 20, rng >> 10, dt * 1e3, st["out_bytes"] / dt / 1e9, st), flush=True)
for thr in (16,):
    for rep in range(2):
        t0 = time.perf_counter(); st = capi.gunzip(p + ".gz", -1, thr); dt = time.perf_counter() - t0
    print("host %d threads: %.1f ms  %.2f GB/s inflated" % (thr, dt * 1e3, st["out_bytes"] / dt / 1e9), flush=True)
subprocess.call("rm -rf %s" % tmp, shell=True)
