"""Parallel gzip inflate of the ingest side (baitstr_b200/csrc/pinflate.cpp) against Python's
gzip/zlib, through the host-only C-ABI hook bstr_gunzip.  The reference opens '*.gz' with the
gzip module (src/fastq.pyx:50-51); whatever that accepts must come out byte-identical here, and
what it rejects (corrupt data, CRC/ISIZE mismatch, truncation) must fail loudly."""
import gzip
import os
import random
import struct
import zlib

import pytest

from tests import util
from baitstr_b200 import capi

CONFIGS = [(4, 1 << 20, 8 << 20), (3, 65536, 1000000), (2, 4096, 4096), (1, 30000, 77777)]


def inflate(tmp_path, z, threads, range_bytes, seg_bytes):
    p, o = str(tmp_path / "x.gz"), str(tmp_path / "x.out")
    open(p, "wb").write(z)
    with open(o, "wb") as fo:
        st = capi.gunzip(p, fo.fileno(), threads, range_bytes, seg_bytes)
    data = open(o, "rb").read()
    assert st["out_bytes"] == len(data) and st["in_bytes"] == len(z)
    return data, st


@pytest.fixture(scope="module")
def fastq():
    return util.synth_fastq(24000, 100, seed=401, tag="z")            # ~5.5 MB


def bgzf(data, block=65280):
    out = []
    for i in range(0, len(data), block):
        blk = data[i:i + block]
        c = zlib.compressobj(6, zlib.DEFLATED, -15)
        d = c.compress(blk) + c.flush()
        out.append(b"\x1f\x8b\x08\x04\0\0\0\0\0\xff\x06\0BC\x02\0" + struct.pack("<H", len(d) + 25) + d +
                   struct.pack("<II", zlib.crc32(blk), len(blk)))
    out.append(bytes.fromhex("1f8b08040000000000ff0600424302001b0003000000000000000000"))      # the empty EOF member
    return b"".join(out)


@pytest.mark.parametrize("level", [1, 6, 9])
def test_single_member_levels(tmp_path, fastq, level):
    z = gzip.compress(fastq, level, mtime=0)
    for cfg in CONFIGS:
        data, st = inflate(tmp_path, z, *cfg)
        assert data == fastq
        assert st["members"] == 1
    # with 64 KiB ranges nearly every range must have been decoded from a start it found itself
    _, st = inflate(tmp_path, z, 3, 65536, 1000000)
    assert st["ranges"] > 20 and st["redone"] <= st["ranges"] // 4


def test_multi_member_bgzf_and_flush_points(tmp_path, fastq):
    parts = [fastq[:1000], fastq[1000:2_000_000], b"", fastq[2_000_000:2_000_100], fastq[2_000_100:]]
    multi = b"".join(gzip.compress(p, lv, mtime=0) for p, lv in zip(parts, (1, 6, 6, 9, 3)))
    c = zlib.compressobj(6, zlib.DEFLATED, 31)
    chunks = []
    for n, i in enumerate(range(0, len(fastq), 300_001)):
        chunks += [c.compress(fastq[i:i + 300_001]), c.flush(zlib.Z_FULL_FLUSH if n % 2 else zlib.Z_SYNC_FLUSH)]
    flushed = b"".join(chunks) + c.flush()
    for z, members in ((multi, 5), (bgzf(fastq), len(fastq) // 65280 + 2), (flushed, 1)):
        assert gzip.decompress(z) == fastq
        for cfg in CONFIGS[:3]:
            data, st = inflate(tmp_path, z, *cfg)
            assert data == fastq and st["members"] == members


def test_block_types_binary_and_repeats(tmp_path, fastq):
    rnd = random.Random(7).randbytes(700_000)
    rep = b"@r\nACGTACGTAC\n+\nIIIIIIIIII\n" * 400_000                    # 10.8 MB from ~20 KB: ratio ~500
    cases = [gzip.compress(fastq[:1_500_000], 0, mtime=0),                # stored blocks only
             gzip.compress(rnd + fastq[:600_000] + rnd[:5000], 6, mtime=0),   # bytes a FASTQ never has: block search must give up cleanly
             gzip.compress(rep, 6, mtime=0),
             gzip.compress(b"", 6, mtime=0),
             gzip.compress(b"@r\nACGT\n+\nIIII\n", 9, mtime=0)]              # one fixed-Huffman block
    for z in cases:
        want = gzip.decompress(z)
        for cfg in CONFIGS[1:3]:
            data, _ = inflate(tmp_path, z, *cfg)
            assert data == want


def test_header_fields_padding_and_trailing_garbage(tmp_path, fastq):
    body = fastq[:2_000_000]
    hdr = b"\x1f\x8b\x08" + bytes([2 | 4 | 8 | 16]) + b"\0\0\0\0\0\x03" + struct.pack("<H", 5) + b"hello" + b"name.fq\0" + b"a comment\0" + b"\x12\x34"
    c = zlib.compressobj(6, zlib.DEFLATED, -15)
    z = hdr + c.compress(body) + c.flush() + struct.pack("<II", zlib.crc32(body), len(body))
    for cfg in CONFIGS[:2]:
        assert inflate(tmp_path, z, *cfg)[0] == body
    # zero padding, then a second member, then more zero padding: accepted (as Python's gzip module does)
    z2 = z + b"\0" * 700 + gzip.compress(fastq[2_000_000:2_500_000], 6, mtime=0)
    assert gzip.decompress(z2 + b"\0" * 33) == fastq[:2_500_000]
    for threads in (3, 0):                                   # the parallel inflater and the serial reader of small files
        assert inflate(tmp_path, z2 + b"\0" * 33, threads, 65536, 1000000)[0] == fastq[:2_500_000]
    # bytes after the last member that are no gzip member: the reference's gzip.open raises
    # ("Not a gzipped file"), so do both readers -- whatever the size of the file
    junk = z2 + b"trailing junk, not gzip" * 9
    with pytest.raises(Exception):
        gzip.decompress(junk)
    for threads in (3, 0):
        p = str(tmp_path / "junk.gz")
        open(p, "wb").write(junk)
        with pytest.raises(capi.BstrError):
            capi.gunzip(p, -1, threads, 65536, 1000000)


def test_serial_reader_matches_parallel(tmp_path, fastq):
    """threads = 0: the zlib-based reader bstr_run_files uses for small files."""
    cases = [gzip.compress(fastq[:300000], 6, mtime=0),
             gzip.compress(fastq[:1000], 1, mtime=0) + gzip.compress(fastq[1000:5000], 9, mtime=0),
             gzip.compress(b"", 6, mtime=0)]
    for z in cases:
        assert inflate(tmp_path, z, 0, 65536, 1000000)[0] == gzip.decompress(z)
    good = gzip.compress(fastq[:300000], 6, mtime=0)
    crc = bytearray(good); crc[-6] ^= 1
    for bad in (good[:len(good) // 2], bytes(crc), fastq[:100000], good[:12]):
        p = str(tmp_path / "bad.gz")
        open(p, "wb").write(bad)
        with pytest.raises(capi.BstrError):
            capi.gunzip(p, -1, 0)


def test_errors_are_loud(tmp_path, fastq):
    good = gzip.compress(fastq, 6, mtime=0)
    flip = bytearray(good); flip[len(flip) // 2] ^= 0x10
    crc = bytearray(good); crc[-6] ^= 1
    isize = bytearray(good); isize[-2] ^= 1
    for bad in (good[:len(good) // 2], bytes(flip), bytes(crc), bytes(isize), fastq[:100000], good[:12], b""):
        p = str(tmp_path / "bad.gz")
        open(p, "wb").write(bad)
        for cfg in CONFIGS[:2]:
            with pytest.raises(capi.BstrError):
                capi.gunzip(p, -1, *cfg)
    with pytest.raises(capi.BstrError):
        capi.gunzip(str(tmp_path / "missing.gz"), -1)
