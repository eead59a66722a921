"""GPU gzip inflate (baitstr_b200/csrc/ginflate.cu, ginflate_dev.cuh) against Python's gzip/zlib,
through the C-ABI hook bstr_gunzip_gpu.  The reference opens '*.gz' with the gzip module
(src/fastq.pyx:50-51): whatever that accepts must come out byte-identical, what it rejects
(corrupt data, CRC/ISIZE mismatch, truncation, bytes that are no gzip member) must fail loudly.
Same corpus as the host inflater's tests (tests/test_inflate_cpu.py), with window and range sizes
small enough that every file spans many windows and ranges."""
import gzip
import struct
import zlib

import pytest

from tests import util
from tests.test_inflate_cpu import bgzf

pytestmark = pytest.mark.gpu

# (window bytes, range bytes): defaults, many ranges per window, many windows
CONFIGS = [(0, 0), (1 << 20, 16 << 10), (128 << 10, 8 << 10)]


@pytest.fixture(scope="module")
def capi():
    from baitstr_b200 import capi as m
    assert m.device_count() > 0, "no CUDA device: the GPU inflater cannot run"
    return m


@pytest.fixture(scope="module")
def fastq():
    return util.synth_fastq(24000, 100, seed=401, tag="z")            # ~5.5 MB


def inflate(capi, tmp_path, z, window, rng):
    p, o = str(tmp_path / "x.gz"), str(tmp_path / "x.out")
    open(p, "wb").write(z)
    with open(o, "wb") as fo:
        st = capi.gunzip_gpu(p, fo.fileno(), 0, window, rng)
    data = open(o, "rb").read()
    assert st["out_bytes"] == len(data) and st["in_bytes"] == len(z)
    return data, st


@pytest.mark.parametrize("level", [1, 6, 9])
def test_single_member_levels(capi, tmp_path, fastq, level):
    z = gzip.compress(fastq, level, mtime=0)
    for cfg in CONFIGS:
        data, st = inflate(capi, tmp_path, z, *cfg)
        assert data == fastq
        assert st["members"] == 1
    # nearly every range must have been decoded from a block start it found itself
    _, st = inflate(capi, tmp_path, z, 1 << 20, 16 << 10)
    assert st["ranges"] > 20 and st["redone"] <= st["ranges"] // 4, st


def test_multi_member_bgzf_and_flush_points(capi, tmp_path, fastq):
    parts = [fastq[:1000], fastq[1000:2_000_000], b"", fastq[2_000_000:2_000_100], fastq[2_000_100:]]
    multi = b"".join(gzip.compress(p, lv, mtime=0) for p, lv in zip(parts, (1, 6, 6, 9, 3)))
    c = zlib.compressobj(6, zlib.DEFLATED, 31)
    flushed = b""
    for i in range(0, len(fastq), 300_000):                            # sync / full flush points: empty stored blocks
        flushed += c.compress(fastq[i:i + 300_000]) + c.flush(zlib.Z_FULL_FLUSH if (i // 300_000) % 2 else zlib.Z_SYNC_FLUSH)
    flushed += c.flush()
    for z in (multi, bgzf(fastq), flushed):
        assert gzip.decompress(z) == fastq
        for cfg in CONFIGS:
            data, _ = inflate(capi, tmp_path, z, *cfg)
            assert data == fastq
    _, st = inflate(capi, tmp_path, bgzf(fastq), 1 << 20, 16 << 10)
    assert st["members"] == len(fastq) // 65280 + 2


def test_stored_fixed_and_tiny(capi, tmp_path, fastq):
    import random
    rnd = random.Random(5)
    noise = bytes(rnd.randrange(256) for _ in range(300_000))          # level 0 and incompressible data: stored blocks
    cases = [gzip.compress(fastq[:700_000], 0, mtime=0), gzip.compress(noise, 6, mtime=0),
             gzip.compress(fastq[:300_000] + noise + fastq[300_000:900_000], 6, mtime=0),
             gzip.compress(b"", 6, mtime=0), gzip.compress(b"A", 6, mtime=0),
             gzip.compress(b"@r\nACGT\n+\nIIII\n", 9, mtime=0)]              # one fixed-Huffman block
    for z in cases:
        want = gzip.decompress(z)
        for cfg in CONFIGS:
            assert inflate(capi, tmp_path, z, *cfg)[0] == want


def test_header_fields_padding_and_trailing_garbage(capi, tmp_path, fastq):
    body = fastq[:2_000_000]
    hdr = b"\x1f\x8b\x08" + bytes([2 | 4 | 8 | 16]) + b"\0\0\0\0\0\x03" + struct.pack("<H", 5) + b"hello" + b"name.fq\0" + b"a comment\0" + b"\x12\x34"
    c = zlib.compressobj(6, zlib.DEFLATED, -15)
    z = hdr + c.compress(body) + c.flush() + struct.pack("<II", zlib.crc32(body), len(body))
    for cfg in CONFIGS:
        assert inflate(capi, tmp_path, z, *cfg)[0] == body
    z2 = z + b"\0" * 700 + gzip.compress(fastq[2_000_000:2_500_000], 6, mtime=0)
    for cfg in CONFIGS:
        assert inflate(capi, tmp_path, z2 + b"\0" * 33, *cfg)[0] == fastq[:2_500_000]
    junk = z2 + b"trailing junk, not gzip" * 9
    p = str(tmp_path / "junk.gz")
    open(p, "wb").write(junk)
    for cfg in CONFIGS:
        with pytest.raises(capi.BstrError):
            capi.gunzip_gpu(p, -1, 0, *cfg)


def test_errors_are_loud(capi, tmp_path, fastq):
    good = gzip.compress(fastq, 6, mtime=0)
    flip = bytearray(good); flip[len(flip) // 2] ^= 0x10
    crc = bytearray(good); crc[-6] ^= 1
    isize = bytearray(good); isize[-2] ^= 1
    for bad in (good[:len(good) // 2], bytes(flip), bytes(crc), bytes(isize), fastq[:100000], good[:12], b""):
        p = str(tmp_path / "bad.gz")
        open(p, "wb").write(bad)
        for cfg in CONFIGS:
            with pytest.raises(capi.BstrError):
                capi.gunzip_gpu(p, -1, 0, *cfg)
    with pytest.raises(capi.BstrError):
        capi.gunzip_gpu(str(tmp_path / "missing.gz"), -1)
    # and a good file still works afterwards
    assert inflate(capi, tmp_path, good, 0, 0)[0] == fastq


def test_larger_file_default_geometry(capi, tmp_path):
    big = util.synth_fastq(400_000, 150, seed=402, tag="big")          # ~130 MB of text
    z = gzip.compress(big, 1, mtime=0)
    data, st = inflate(capi, tmp_path, z, 8 << 20, 0)                   # several windows
    assert data == big and st["ranges"] > 300 and st["redone"] <= 8, st
