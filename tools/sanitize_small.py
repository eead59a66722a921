#!/usr/bin/env python3
"""Small inputs through the round-2 device code, meant to run under compute-sanitizer:
   compute-sanitizer --tool memcheck python tools/sanitize_small.py"""
import gzip, os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.update(BSTR_GZ_GPU_MIN="0", BSTR_GZ_WINDOW=str(256 << 10), BSTR_GZ_RANGE=str(16 << 10))
from baitstr_b200 import capi
from tests import util
from tests.test_inflate_cpu import bgzf
raw = util.synth_fastq(6000, 100, seed=5, p_str=0.5)
tmp = tempfile.mkdtemp(prefix="san_")
cases = {"a.fq.gz": gzip.compress(raw, 1), "b.fq.gz": bgzf(raw), "c.fq.gz": gzip.compress(raw[:400000], 9) + gzip.compress(raw[400000:-1], 0),
         "d.fq.gz": gzip.compress(b"@r\nACGT\n+\nIIII\n", 9)}
for name, z in cases.items():
    p = os.path.join(tmp, name)
    open(p, "wb").write(z)
    for win, rng in ((0, 0), (128 << 10, 8 << 10)):
        st = capi.gunzip_gpu(p, -1, 0, win, rng)
        assert st["out_bytes"] == len(gzip.decompress(z)), (name, st)
with capi.Context() as ctx, open(os.devnull, "wb") as dn:
    want = ctx.scan_batch(raw)[0]
    outp = os.path.join(tmp, "o.fq")
    with open(outp, "wb") as fo:
        ctx.run_files([os.path.join(tmp, "a.fq.gz"), os.path.join(tmp, "b.fq.gz")], out_fd=fo.fileno(), err_fd=dn.fileno())
    assert open(outp, "rb").read() == want + want
    assert ctx.scan_batch_noalloc(raw)[0] == want
    assert ctx.scan_batch_noalloc(raw[:100000], final=False)[1]["consumed"] < 100000
bad = bytearray(cases["a.fq.gz"]); bad[len(bad) // 2] ^= 4
p = os.path.join(tmp, "bad.fq.gz"); open(p, "wb").write(bytes(bad[:len(bad) * 3 // 4]))
try:
    capi.gunzip_gpu(p, -1, 0, 128 << 10, 8 << 10)
except capi.BstrError as e:
    print("corrupt input rejected:", str(e)[:80])
print("sanitize_small: ok")
