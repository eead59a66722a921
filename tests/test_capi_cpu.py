"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/baitstr_select.h declares, and refuses to run without a CUDA device (no fallback).
No compute calls here."""
import ctypes
import os
import re
import subprocess
import sys

import pytest

from tests import util

ROOT = util.ROOT


def test_header_symbols_exported():
    from baitstr_b200 import capi
    lib = capi.load_library()
    hdr = open(os.path.join(ROOT, "include", "baitstr_select.h")).read()
    declared = set(re.findall(r"\b(bstr_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(capi.EXPORTS), declared ^ set(capi.EXPORTS)
    for name in declared:
        assert getattr(lib, name) is not None


def test_struct_sizes_match_header():
    """ctypes mirrors vs the C compiler's view of the header."""
    from baitstr_b200 import capi
    src = '#include "baitstr_select.h"\n#include <stdio.h>\nint main(){printf("%zu %zu %zu %zu %zu\\n",' \
          'sizeof(bstr_params),sizeof(bstr_batch_out),sizeof(bstr_stats),sizeof(bstr_bench_out),' \
          'sizeof(bstr_read_result));return 0;}'
    exe = "/tmp/bstr_sizes_%d" % os.getpid()
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), "-x", "c", "-", "-o", exe],
                   input=src.encode(), check=True)
    out = subprocess.check_output([exe]).split()
    os.unlink(exe)
    assert [int(x) for x in out] == [ctypes.sizeof(capi.Params), ctypes.sizeof(capi.BatchOut),
                                     ctypes.sizeof(capi.Stats), ctypes.sizeof(capi.BenchOut), 16]


def test_no_device_is_a_hard_error():
    from baitstr_b200 import capi
    if capi.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(capi.BstrError):
        capi.Context()


def test_product_does_not_touch_oracle():
    """The product package and its C sources never reference oracle/ or tests/."""
    pkg = os.path.join(ROOT, "baitstr_b200")
    for dp, dn, fn in os.walk(pkg):
        for f in fn:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".inc", ".c")):
                text = open(os.path.join(dp, f), errors="replace").read()
                assert "import oracle" not in text and "from oracle" not in text, f
                assert "str_oracle" not in text and "libemul" not in text, f


def _cli(*args):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bin", "select_STR_reads")] + list(args),
                          capture_output=True, text=True)


def test_cli_contract_without_gpu():
    """Exit codes / streams of /root/reference/src/select_STR_reads:75-78,93-95,124-132."""
    r = _cli("-h")
    assert r.returncode == 0 and "usage:" in r.stderr and r.stdout == ""
    # the usage text is the reference's, byte for byte (/root/reference/src/select_STR_reads:2-46, version filled in)
    import hashlib
    assert hashlib.md5(r.stderr.encode()).hexdigest() == "27021ebfa48636253a26e216e44e9f19"
    if os.path.exists("/root/reference/src/select_STR_reads"):
        ref = open("/root/reference/src/select_STR_reads").read()
        a = ref.index('"""') + 3
        assert r.stderr == ref[a:ref.index('"""', a)] % {"scriptVersion": "0.3.1"} + "\n"
    r = _cli("-v")
    assert r.returncode == 0 and r.stderr.splitlines()[0] == "Program: select_STR_reads"
    assert r.stderr.splitlines()[1] == "Version: 0.3.1"
    r = _cli("--bogus")
    assert r.returncode == 2 and "option --bogus not recognized" in r.stdout and "usage:" in r.stderr
    r = _cli()
    assert r.returncode == 3 and "usage:" in r.stderr
    r = _cli("--flank", "5")          # unique-prefix long option, then no files
    assert r.returncode == 3


def test_tile_geometry_invariants():
    """pick_geometry (csrc/select_host.cu) over record sizes and kernel instantiations: what the
    fused kernel relies on (fused.cuh) must hold for every pick.  No GPU involved."""
    import ctypes as C
    from baitstr_b200 import capi
    lib = capi.load_library()
    lib.bstr_geometry_probe.argtypes = [C.c_double, C.c_int, C.POINTER(C.c_uint32)]
    lib.bstr_geometry_probe.restype = C.c_int
    out = (C.c_uint32 * 8)()
    picks = {}
    for w in (2, 4, 6, 8):
        for rec in [16, 40, 90, 124, 223, 224, 260, 323, 400, 524, 700, 1024, 1500, 2100, 4000, 9000]:
            assert lib.bstr_geometry_probe(float(rec), w, out) == 0
            tile, thr, region, chunks, cpt, recmax, stage, smem = list(out)
            ctx = (w, rec, list(out))
            assert tile % 16 == 0 and tile >= 8192 and region == tile + 4096 and region <= 65504, ctx
            assert thr % 32 == 0 and 64 <= thr <= (192 if w >= 6 else 256), ctx
            assert chunks == region // 16 and cpt * thr >= chunks and cpt <= 16, ctx     # FT_MAXCPT
            assert recmax == 2 * thr and smem <= 200 * 1024, ctx                           # FT_MAXSMEM
            if rec >= 90:
                assert 2 * (smem + 1024) <= 228 * 1024, ctx                                # two CTAs per SM
                assert stage >= tile * (1 + 48.0 / rec), ctx
            picks[(w, rec)] = (tile, thr)
    # the picks the round-1 measurements were taken with
    assert picks[(2, 223)] == (41984, 256) and picks[(4, 323)] == (40928, 192) and picks[(6, 524)] == (45056, 192)
    assert lib.bstr_geometry_probe(0.0, 2, out) != 0
