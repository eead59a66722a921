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
def test_gpu_counts_the_batch_the_select_stage_left_resident(capi):
    """bstr_kmers_add_resident: the reads cross PCIe once for the select stage and the counting pass"""
    raws = [genome_reads(30000, 150, 200000, seed=5), genome_reads(20000, 100, 200000, seed=6)]
    with capi.Context() as ctx:
        want = gpu_pairs(ctx, raws, 27)[0]
        with ctx.kmer_counter(27, 1 << 22) as kc:
            with pytest.raises(capi.BstrError) as ei:          # nothing scanned yet on this context
                kc.add_resident(1000)
            assert "resident" in str(ei.value)
            selected = b""
            for raw in raws:
                cut = raw + raw[:200]                          # a batch that ends inside a record: only whole records count
                out, info = ctx.scan_batch(cut, final=False)
                assert info["consumed"] == len(raw)
                kc.add_resident(info["consumed"])
                selected += out
            with pytest.raises(capi.BstrError):                # more than the batch holds
                kc.add_resident(len(cut) + 1)
            kmers, counts = kc.finish()
            assert [(int(k), int(c)) for k, c in zip(kmers, counts)] == want
            assert kc.stats()["records"] == 50000
        from oracle import oracle
        assert selected == b"".join(oracle.select_buffer(r) for r in raws)     # and the select output is what it always is


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


# ---------------------------------------------------------------- the extension walk (bstr_kmers_extend)
#
# The rest of extend_STR_reads over the counted table: extend_STR_reads.c:195-998.  Checkers: the
# reference's own binary (oracle/_ref/downstream/extend_STR_reads, compiled in place) and the golden it
# wrote for the README's walk-through (tests/golden/extend_readme.txt, make_extend_golden.sh).
# not gpu: the product's walk (csrc/extend_core.cuh, extend_plan.h) built for the host with a sequential
#          executor (tests/emul/emul_extend.cpp) over the table the reference's counting function dumps
# gpu:     counting + walks on the device through the C ABI, and through bin/extend_STR_reads

EXTEND_BIN = os.path.join(ROOT, "oracle", "_ref", "downstream", "extend_STR_reads")
MERGE_BIN = os.path.join(ROOT, "oracle", "_ref", "downstream", "merge_STR_reads")
EMUL_SRC = os.path.join(ROOT, "tests", "emul", "emul_extend.cpp")
EMUL_BIN = os.path.join(ROOT, "tests", "emul", "_build", "emul_extend")
EXTEND_USAGE_MD5 = "6d81c3d9b46af6dc328c63bd14a88b84"     # md5 of `extend_STR_reads --help` of the reference (0.3.1), stdout
needs_bins = pytest.mark.skipif(not (os.path.exists(EXTEND_BIN) and os.path.exists(MERGE_BIN) and os.path.exists(REF_BIN)),
                                reason="oracle/_ref/downstream not built (oracle/build_downstream.sh)")

DIPLOID = [
    # name, (genome_len, nloci, coverage, read length, seed, repeat_len), [(klen, flanks, min, max)]
    ("150bp", (40000, 30, 40, 150, 1, 0), [(27, 1024, 2, 255), (27, 100, 2, 255), (21, 1024, 4, 60)]),
    ("100bp_with_repeats", (50000, 40, 30, 100, 2, 400), [(27, 1024, 2, 255), (31, 1024, 2, 255), (27, 1024, 10, 255)]),
    ("250bp", (30000, 12, 50, 250, 3, 0), [(27, 1024, 2, 255), (25, 300, 3, 255)]),
]


def emul_extend_bin():
    csrc = os.path.join(ROOT, "baitstr_b200", "csrc")
    deps = [EMUL_SRC, os.path.join(csrc, "extend_core.cuh"), os.path.join(csrc, "extend_plan.h")]
    if not os.path.exists(EMUL_BIN) or os.path.getmtime(EMUL_BIN) < max(os.path.getmtime(d) for d in deps):
        os.makedirs(os.path.dirname(EMUL_BIN), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-Wall", "-o", EMUL_BIN, EMUL_SRC])
    return EMUL_BIN


_diploid = {}


def diploid_case(name, tmp_path):
    """reads -> select (oracle) -> merge (the reference's binary): paths of blocks and the two read files"""
    if name not in _diploid:
        from oracle import oracle
        args = dict((n, a) for n, a, _ in DIPLOID)[name]
        raw = util.diploid_str_reads(args[0], args[1], args[2], args[3], args[4], repeat_len=args[5])
        recs = raw.split(b"\n@")
        half = len(recs) // 2
        r1, r2 = b"\n@".join(recs[:half]) + b"\n", b"@" + b"\n@".join(recs[half:])
        sel = oracle.select_buffer(raw, n=3)
        p = tmp_path / "sel.fq"
        p.write_bytes(sel)
        merged = subprocess.run([MERGE_BIN, "--all", "--min_threshold=3", "27", str(p)], check=True, capture_output=True).stdout
        _diploid[name] = (merged, r1, r2)
    merged, r1, r2 = _diploid[name]
    paths = []
    for fn, data in (("merged.fq", merged), ("r1.fq", r1), ("r2.fq", r2)):
        q = tmp_path / fn
        q.write_bytes(data)
        paths.append(str(q))
    return paths, (merged, r1, r2)


def reference_extend(paths, klen, flanks, lo, hi):
    return subprocess.run([EXTEND_BIN, "--min_threshold=%d" % lo, "--max_threshold=%d" % hi, "--flanks=%d" % flanks, "4000000", "20", str(klen)] + paths,
                          check=True, capture_output=True).stdout


def test_extend_golden_is_the_readme_known_answer():
    """README.md:265-287: one contig for the shipped test_data"""
    with open(os.path.join(ROOT, "tests", "golden", "extend_readme.txt"), "rb") as f:
        assert f.read().split(b"\n")[0] == b">Block1\tTC:7,5:191:205"


@needs_bins
@pytest.mark.parametrize("case", DIPLOID, ids=[d[0] for d in DIPLOID])
def test_host_build_of_the_walk_matches_the_reference_binary(case, tmp_path):
    name, _, runs = case
    paths, _ = diploid_case(name, tmp_path)
    contigs = 0
    for klen, flanks, lo, hi in runs:
        want = reference_extend(paths, klen, flanks, lo, hi)
        dump = tmp_path / "kmers.txt"
        dump.write_bytes(subprocess.run([REF_BIN, str(klen), str(lo), str(hi), "4000000"] + paths, check=True, capture_output=True).stdout)
        got = subprocess.run([emul_extend_bin(), str(klen), str(flanks), str(dump), paths[0]], check=True, capture_output=True).stdout
        assert got == want, (klen, flanks, lo, hi)
        contigs += want.count(b">")
        if flanks >= 1024:
            assert max(len(ln) for ln in want.split(b"\n")) > 800     # the walks go far beyond the blocks
    assert contigs > 100


def test_extend_cli_usage_is_the_reference_text():
    import hashlib
    import sys
    cli = os.path.join(ROOT, "bin", "extend_STR_reads")
    r = subprocess.run([sys.executable, cli, "--help"], capture_output=True)
    assert r.returncode == 0 and hashlib.md5(r.stdout).hexdigest() == EXTEND_USAGE_MD5
    if os.path.exists(EXTEND_BIN):
        assert subprocess.run([EXTEND_BIN, "--help"], capture_output=True).stdout == r.stdout
    r = subprocess.run([sys.executable, cli, "4000", "20", "27", "blocks.fq", "reads1.fq"], capture_output=True)   # argc < 7 (:1046-1049)
    assert r.returncode == 1 and hashlib.md5(r.stdout).hexdigest() == EXTEND_USAGE_MD5
    r = subprocess.run([sys.executable, cli, "--flank=3", "4000", "20", "27", "a", "b", "c"], capture_output=True)
    assert r.returncode == 1 and b"unknown option flank" in r.stderr


@pytest.mark.gpu
def test_gpu_extend_test_data_matches_the_reference_golden(capi):
    blocks = open(os.path.join(ROOT, "tests", "golden", "merge_readme_k27.txt"), "rb").read()
    reads = [util.gz_read("testdata_%d.fq.gz" % i) for i in (1, 2)]
    with capi.Context() as ctx:
        out, st = ctx.extend_reads(blocks, reads, 27, expected_kmers=200000)
    assert out == open(os.path.join(ROOT, "tests", "golden", "extend_readme.txt"), "rb").read()
    assert st["blocks"] == 1 and st["contigs"] == 1 and st["ends_not_extended"] == 0 and st["kmers_kept"] >= 2013


@pytest.mark.gpu
@pytest.mark.parametrize("case", DIPLOID, ids=[d[0] for d in DIPLOID])
def test_gpu_extend_matches_the_reference_binary(capi, case, tmp_path):
    if not (os.path.exists(EXTEND_BIN) and os.path.exists(MERGE_BIN)):
        pytest.skip("oracle/_ref/downstream not built")
    name, _, runs = case
    paths, (merged, r1, r2) = diploid_case(name, tmp_path)
    with capi.Context() as ctx:
        for klen, flanks, lo, hi in runs:
            out, st = ctx.extend_reads(merged, [r1, r2], klen, lo, hi, flanks, expected_kmers=1 << 22)
            assert out == reference_extend(paths, klen, flanks, lo, hi), (klen, flanks, lo, hi)
            assert st["blocks"] == merged.count(b"\n") // 4 and st["contigs"] == out.count(b">")


@pytest.mark.gpu
def test_gpu_extend_cli_and_errors(capi, tmp_path):
    import sys
    cli = os.path.join(ROOT, "bin", "extend_STR_reads")
    paths = []
    for fn, data in (("merged.reads.str.fq", open(os.path.join(ROOT, "tests", "golden", "merge_readme_k27.txt"), "rb").read()),
                     ("reads_1.fq", util.gz_read("testdata_1.fq.gz")), ("reads_2.fq.gz", gzip.compress(util.gz_read("testdata_2.fq.gz")))):
        p = tmp_path / fn
        p.write_bytes(data)
        paths.append(str(p))
    r = subprocess.run([sys.executable, cli, "4000", "20", "27"] + paths, capture_output=True)       # test_data/Makefile:10-12
    assert r.returncode == 0 and r.stdout == open(os.path.join(ROOT, "tests", "golden", "extend_readme.txt"), "rb").read(), r.stderr[-2000:]
    assert b"Expecting 23633 kmers in this dataset with haploid genome size 4000 bps." in r.stderr
    r = subprocess.run([sys.executable, cli, "--flanks=60", "4000", "20", "28"] + paths, capture_output=True)
    assert r.returncode == 0 and b"using 27" in r.stderr
    if os.path.exists(EXTEND_BIN):
        plain = tmp_path / "reads_2.fq"
        plain.write_bytes(util.gz_read("testdata_2.fq.gz"))
        assert r.stdout == reference_extend(paths[:2] + [str(plain)], 27, 60, 2, 255)
    blocks = open(paths[0], "rb").read()
    with capi.Context() as ctx:
        with pytest.raises(capi.BstrError) as ei:
            ctx.extend_reads(blocks, [util.gz_read("testdata_1.fq.gz")], 27, min_threshold=1)
        assert "min_threshold" in str(ei.value)
        with pytest.raises(capi.BstrError) as ei:
            ctx.extend_reads(blocks, [util.gz_read("testdata_1.fq.gz")], 27, flanks=5000)
        assert "flanks" in str(ei.value)
        with pytest.raises(capi.BstrError) as ei:            # not the merge stage's output
            ctx.extend_reads(util.gz_read("testdata_1.fq.gz")[:2000].rsplit(b"\n@", 1)[0] + b"\n", [], 27)
        assert "parsing the name" in str(ei.value)
        out, st = ctx.extend_reads(blocks, [util.gz_read("testdata_1.fq.gz")], 31)      # klen > the block's left flank? no: 60 >= 31; still one contig or none, and no crash
        assert st["blocks"] == 1
