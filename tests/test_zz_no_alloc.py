"""bstr_scan_batch_noalloc (include/baitstr_select.h): the fused kernel with all of its device
memory in one static array, for callers that may not allocate.  No counterpart in the
reference.

Checked through the C ABI against the existing entry point bstr_scan_batch (itself byte-identical
to the oracle and the reference's goldens in tests/test_gpu_parity.py): the bar is the one every
test here uses for byte work -- bit-exact output and equal counts.  Calls are made back to back on
purpose: the array is shared and keeps the previous call's contents.

This file sorts last (pytest --collect-only) so that a failure here hides no other test under -x,
and it passes alone in a fresh process as well as after the other tests (it relies on no state of
theirs: its own contexts, its own inputs)."""
import pytest

from tests import util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def capi():
    from baitstr_b200 import capi as m
    assert m.device_count() > 0, "no CUDA device: the product path cannot run (no fallback)"
    return m


def both(ctx, raw, final=True):
    want, winfo = ctx.scan_batch(raw, final=final)
    got, ginfo = ctx.scan_batch_noalloc(raw, final=final)
    assert got == want
    assert ginfo == winfo
    return want


def test_two_calls_in_a_row_match_the_allocating_entry_point(capi):
    big = util.synth_fastq(9000, 100, seed=811, p_str=0.5)              # ~2 MB, many tiles
    small = util.synth_fastq(300, 150, seed=812, p_str=0.7, tag="s")    # fewer tiles than the call before: stale look-back words
    lower = util.synth_fastq(1200, 76, seed=813, p_str=0.4, tag="l").lower().replace(b"@l", b"@L")
    with capi.Context() as ctx:
        a = both(ctx, big)
        b = both(ctx, small)                       # second call in a row, smaller than the first
        assert both(ctx, big) == a                 # and the first input again: same bytes as before
        assert both(ctx, small) == b
        assert len(a) > 0 and len(b) > 0 and a != b
        both(ctx, lower)
        both(ctx, big[:-1])                        # unterminated last line (fastq.pyx:90)
        both(ctx, b"")                             # empty batch
        # not final: only whole records are consumed, as with bstr_scan_batch
        cut = big[:len(big) // 2 + 37]
        want, winfo = ctx.scan_batch(cut, final=False)
        got, ginfo = ctx.scan_batch_noalloc(cut, final=False)
        assert got == want and ginfo == winfo and ginfo["consumed"] < len(cut)
    # other flag sets, a fresh context each (the array outlives the contexts)
    for kw in (dict(n=3, rmhm=True), dict(ks=[2, 3], f=41), dict(altalgo=False, f=10), dict(illumina=True, n=3, f=29)):
        with capi.Context(**kw) as ctx:
            both(ctx, util.synth_fastq(2500, 150, seed=820, p_str=0.6, tag="f"))
            both(ctx, small)
    # the shipped test_data: the reference's golden through the no-allocation path
    import hashlib
    with capi.Context() as ctx:
        out = b"".join(ctx.scan_batch_noalloc(util.gz_read("testdata_%d.fq.gz" % i))[0] for i in (1, 2))
    assert hashlib.md5(out).hexdigest() == util.manifest()["testdata_default"]["md5"]


def test_limits_and_errors_are_loud(capi):
    rec = b"@r\n" + b"ACGT" * 25 + b"\n+\n" + b"I" * 100 + b"\n"
    with capi.Context() as ctx:
        # beyond the stated maximum: refused with a clear message, nothing is run
        too_big = rec * (capi.NOALLOC_MAX_BYTES // len(rec) + 1)
        assert len(too_big) > capi.NOALLOC_MAX_BYTES
        with pytest.raises(capi.BstrError) as ei:
            ctx.scan_batch_noalloc(too_big)
        assert ei.value.code == -14 and "maximum" in str(ei.value) and str(capi.NOALLOC_MAX_BYTES) in str(ei.value)
        # exactly at the maximum it runs, and matches
        at_max = (rec * (capi.NOALLOC_MAX_BYTES // len(rec)))
        both(ctx, at_max)
        # a caller buffer that is too small: error, and the size needed is reported
        raw = util.synth_fastq(2000, 100, seed=830, p_str=0.9)
        want, _ = ctx.scan_batch(raw)
        with pytest.raises(capi.BstrError) as ei:
            ctx.scan_batch_noalloc(raw, out_capacity=len(want) - 1)
        assert ei.value.code == -14 and ei.value.info["out_bytes"] == len(want)
        assert ctx.scan_batch_noalloc(raw, out_capacity=len(want))[0] == want
        # malformed FASTQ: same code and record index as the allocating entry point
        lines = raw.split(b"\n")
        lines[4 * 1234 + 2] = b"-"
        bad = b"\n".join(lines)
        with pytest.raises(capi.BstrError) as e1:
            ctx.scan_batch(bad)
        with pytest.raises(capi.BstrError) as e2:
            ctx.scan_batch_noalloc(bad)
        assert (e2.value.code, e2.value.bad_record) == (e1.value.code, e1.value.bad_record) == (-12, 1234)
        # an input the fused kernel's geometry cannot hold (a record longer than a tile's whole window,
        # in the middle of the batch) is refused here, not re-run through the generic kernels (they
        # need allocated tables); a long record that still fits a window is simply processed
        fits = b"@long\n" + b"ACGT" * 500 + b"\n+\n" + b"I" * 2000 + b"\n"
        both(ctx, raw + fits + raw)
        long_rec = b"@long\n" + b"ACGT" * 12000 + b"\n+\n" + b"I" * 48000 + b"\n"
        assert ctx.scan_batch(raw + long_rec + raw)[1]["n_records"] == 4001
        with pytest.raises(capi.BstrError) as ei:
            ctx.scan_batch_noalloc(raw + long_rec + raw)
        assert ei.value.code == -14 and "bstr_scan_batch" in str(ei.value)
        # and the array is fine afterwards
        assert ctx.scan_batch_noalloc(raw)[0] == want
