"""bstr_kmers_* (include/baitstr_select.h): the k-mer counting pass of extend_STR_reads
(/root/reference/src/extend_STR_reads.c:31-194) on the GPU -- SURVEY.md 8(f) rank 3, first slice.

The bar is the one for all integer work here: identical (k-mer, count) sets.  Three checkers:
  * tests/golden/kmers_testdata_k27.txt.gz -- written by the reference's own function for the
    shipped test_data (tests/golden/make_kmer_golden.sh);
  * oracle/_ref/downstream/kmer_count_ref -- that function compiled from the reference's sources
    in place (oracle/build_downstream.sh); it travels to the GPU box with the snapshot;
  * oracle/kmer_count.py -- the CPU restatement, pinned here against both of the above.
"""
import gzip
import os
import random
import subprocess

import numpy as np
import pytest

from oracle import kmer_count
from tests import util
from tests.util import genome_reads

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "downstream", "kmer_count_ref")
GOLDEN = os.path.join(ROOT, "tests", "golden", "kmers_testdata_k27.txt.gz")


def golden_pairs():
    with gzip.open(GOLDEN, "rt") as f:
        return [tuple(int(x) for x in line.split()) for line in f]


def reference_pairs(tmp_path, raws, klen, lo, hi, expected=200000):
    files = []
    for i, raw in enumerate(raws):
        p = tmp_path / ("reads_%d_%d.fq" % (klen, i))
        p.write_bytes(raw)
        files.append(str(p))
    out = subprocess.run([REF_BIN, str(klen), str(lo), str(hi), str(expected)] + files, check=True,
                         capture_output=True).stdout
    return [tuple(int(x) for x in line.split()) for line in out.decode().splitlines()]


def random_reads(n, lo, hi, alphabet, seed, tag=b"v"):
    rng = random.Random(seed)
    out = []
    for i in range(n):
        ln = rng.randint(lo, hi)
        out.append(b"@%s%d\n%s\n+\n%s\n" % (tag, i, "".join(rng.choice(alphabet) for _ in range(ln)).encode(), b"I" * ln))
    return b"".join(out)


CASES = [
    # name, batches, klen, min, max
    ("plain", lambda: [util.synth_fastq(2000, 100, seed=1, p_n=0, p_lower=0)], 21, 2, 255),
    ("n_and_lower_case", lambda: [util.synth_fastq(2000, 100, seed=2, p_n=0.3, p_lower=0.2)], 21, 2, 255),
    ("saturates_at_255", lambda: [util.synth_fastq(3000, 80, seed=3, p_n=0.1)], 5, 2, 255),
    ("saturated_ones_cut_by_max", lambda: [util.synth_fastq(3000, 80, seed=3, p_n=0.1)], 5, 2, 254),
    ("longest_k", lambda: [util.synth_fastq(3000, 120, seed=4, p_n=0.1, p_str=0.8)], 31, 2, 255),
    ("shortest_k", lambda: [util.synth_fastq(300, 50, seed=5, p_n=0.1)], 2, 2, 255),
    ("two_files_thresholds", lambda: [util.synth_fastq(1500, 100, seed=6, p_str=0.9),
                                      util.synth_fastq(1500, 100, seed=6, p_str=0.9, tag="u")], 27, 3, 20),
    ("reads_shorter_than_k", lambda: [random_reads(3000, 1, 60, "ACGTN", 9)], 17, 2, 255),
    ("bytes_outside_the_alphabet", lambda: [util.synth_fastq(2000, 100, seed=7, p_dot=0.5)], 11, 2, 255),
    ("iupac_and_digits", lambda: [random_reads(1500, 30, 90, "ACGTRYKMSWnacgt.-*0", 10)], 13, 2, 255),
]


# ---------------------------------------------------------------- CPU: the oracle is pinned

def test_golden_is_the_reference_defaults_output():
    pairs = golden_pairs()
    assert len(pairs) == 2013 and pairs == sorted(pairs)
    assert all(2 <= c <= 255 for _, c in pairs)


def test_restatement_matches_the_golden():
    raws = [util.gz_read("testdata_%d.fq.gz" % i) for i in (1, 2)]
    assert kmer_count.count_kmers(raws, 27) == golden_pairs()


@pytest.mark.skipif(not os.path.exists(REF_BIN), reason="oracle/_ref/downstream not built (oracle/build_downstream.sh)")
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_restatement_matches_the_reference_binary(case, tmp_path):
    _, make, klen, lo, hi = case
    raws = make()
    assert kmer_count.count_kmers(raws, klen, lo, hi) == reference_pairs(tmp_path, raws, klen, lo, hi)


@pytest.mark.skipif(not os.path.exists(REF_BIN), reason="oracle/_ref/downstream not built (oracle/build_downstream.sh)")
def test_reference_binary_reproduces_the_golden(tmp_path):
    raws = [util.gz_read("testdata_%d.fq.gz" % i) for i in (1, 2)]
    assert reference_pairs(tmp_path, raws, 27, 2, 255, 100000) == golden_pairs()


# ---------------------------------------------------------------- GPU: parity through the C ABI

@pytest.fixture(scope="module")
def capi():
    from baitstr_b200 import capi as m
    assert m.device_count() > 0, "no CUDA device: the product path cannot run (no fallback)"
    return m


def gpu_pairs(ctx, raws, klen, lo=2, hi=255, expected=1 << 20):
    kmers, counts, st = ctx.count_kmers(raws, klen, lo, hi, expected_kmers=expected)
    assert kmers.dtype == np.uint64 and counts.dtype == np.uint8 and len(kmers) == len(counts)
    return [(int(k), int(c)) for k, c in zip(kmers, counts)], st


@pytest.mark.gpu
def test_gpu_test_data_matches_the_reference_golden(capi):
    raws = [util.gz_read("testdata_%d.fq.gz" % i) for i in (1, 2)]
    with capi.Context() as ctx:
        got, st = gpu_pairs(ctx, raws, 27)
        assert got == golden_pairs()
        nrec = sum(r.count(b"\n") for r in raws) // 4
        assert st["records"] == nrec and st["distinct"] >= len(got) and st["kmers"] > st["distinct"]
        # one batch or many: the same table
        whole = b"".join(raws)
        recs = whole.split(b"\n")
        cut = 4 * 777
        parts = [b"\n".join(recs[:cut]) + b"\n", b"\n".join(recs[cut:])]
        assert gpu_pairs(ctx, parts, 27)[0] == got
        assert gpu_pairs(ctx, [whole], 27, 3, 9)[0] == [(k, c) for k, c in got if 3 <= c <= 9]


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_gpu_matches_the_reference_binary(capi, case, tmp_path):
    _, make, klen, lo, hi = case
    raws = make()
    want = kmer_count.count_kmers(raws, klen, lo, hi)
    if os.path.exists(REF_BIN):
        assert reference_pairs(tmp_path, raws, klen, lo, hi) == want
    with capi.Context() as ctx:
        got, st = gpu_pairs(ctx, raws, klen, lo, hi)
    assert got == want
    assert st["slots"] >= 2 * st["distinct"]


@pytest.mark.gpu
def test_gpu_larger_input_against_the_reference_binary(capi, tmp_path):
    if not os.path.exists(REF_BIN):
        pytest.skip("oracle/_ref/downstream not built")
    raw = genome_reads(200000, 150, 500000, seed=77)
    with capi.Context() as ctx:
        got, st = gpu_pairs(ctx, [raw], 27, expected=1 << 25)
    assert st["records"] == 200000 and st["kmers"] == 200000 * (150 - 27 + 1)
    assert got == reference_pairs(tmp_path, [raw], 27, 2, 255, expected=30000000)


@pytest.mark.gpu
def test_gpu_errors_are_loud(capi):
    raw = util.synth_fastq(4000, 100, seed=11)
    with capi.Context() as ctx:
        with pytest.raises(capi.BstrError) as ei:          # the Bloom filter's false positives would show
            ctx.count_kmers([raw], 21, min_threshold=1)
        assert "min_threshold" in str(ei.value)
        with pytest.raises(capi.BstrError) as ei:          # kmer.c:11 asserts klen < 32
            ctx.count_kmers([raw], 32)
        assert "klen" in str(ei.value)
        with pytest.raises(capi.BstrError) as ei:          # 4000 * 80 k-mers do not fit 2048 slots
            ctx.count_kmers([raw], 21, expected_kmers=1000)
        assert "full" in str(ei.value)
        with pytest.raises(capi.BstrError) as ei:          # half a record
            ctx.count_kmers([b"\n".join(raw.split(b"\n")[:42]) + b"\n"], 21)
        assert "4-line records" in str(ei.value) or "whole" in str(ei.value)
        with pytest.raises(capi.BstrError):                # no final newline
            ctx.count_kmers([raw[:-1]], 21)
        assert ctx.count_kmers([], 21)[0].size == 0
        assert ctx.count_kmers([raw], 21)[2]["records"] == 4000   # the context is fine afterwards
