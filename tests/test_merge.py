"""bstr_merge_reads (include/baitstr_select.h): merge_STR_reads on the GPU -- SURVEY.md 8(f) rank 2,
first slice.  Replaces MergeShortTandemRepeatReads, /root/reference/src/merge_STR_reads.c:308-509.

The bar: the bytes the reference's own binary prints -- same blocks, same consensus sequences and
qualities, same order and numbering.  Checkers:
  * tests/golden/merge_*.txt -- written by the reference's merge_STR_reads for the
    reference-generated select goldens of the shipped test_data (tests/golden/make_merge_golden.sh);
  * oracle/_ref/downstream/merge_STR_reads -- the reference's source compiled in place
    (oracle/build_downstream.sh); it travels to the GPU box with the snapshot.
not gpu: the product's host sources (merge_plan.h: parsing, keys, chains, the replayed table order,
         printing; merge_core.cuh / merge_step.cuh: the per-read step) built for the host with a
         sequential executor (tests/emul/emul_merge.cpp) against both checkers
gpu:     the same through the C ABI, where a warp walks each chain
"""
import os
import subprocess

import pytest

from oracle import oracle
from tests import util

ROOT = util.ROOT
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "downstream", "merge_STR_reads")
EMUL_SRC = os.path.join(ROOT, "tests", "emul", "emul_merge.cpp")
EMUL_BIN = os.path.join(ROOT, "tests", "emul", "_build", "emul_merge")
CSRC = os.path.join(ROOT, "baitstr_b200", "csrc")

GOLDENS = [
    # golden file, select golden it was made from, klength, min, max, all
    ("merge_readme_k27.txt", "testdata_readme.out.gz", 27, 4, 10000, False),
    ("merge_readme_k27_all_min2.txt", "testdata_readme.out.gz", 27, 2, 10000, True),
    ("merge_default_k27_all_min2.txt", "testdata_default.out.gz", 27, 2, 10000, True),
    ("merge_default_k23_min2_max6.txt", "testdata_default.out.gz", 23, 2, 6, False),
]

SYNTH = [
    # name, (nloci, coverage, read length, seed, shared_flank), [(klength, min, max, all)]
    ("150bp", (150, 40, 150, 1, 0.0), [(27, 4, 10000, False), (27, 2, 10000, True), (21, 2, 10000, True)]),
    ("100bp_shared_flanks", (250, 25, 100, 2, 0.2), [(27, 2, 10000, True), (31, 3, 50, True), (26, 2, 10000, False)]),
    ("250bp_long_flanks", (100, 60, 250, 3, 0.1), [(27, 4, 10000, False), (21, 2, 10000, True)]),
]


def golden(name):
    with open(os.path.join(ROOT, "tests", "golden", name), "rb") as f:
        return f.read()


def emul_bin():
    deps = [EMUL_SRC] + [os.path.join(CSRC, h) for h in ("merge_plan.h", "merge_core.cuh", "merge_step.cuh")]
    if not os.path.exists(EMUL_BIN) or os.path.getmtime(EMUL_BIN) < max(os.path.getmtime(d) for d in deps):
        os.makedirs(os.path.dirname(EMUL_BIN), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-Wall", "-o", EMUL_BIN, EMUL_SRC])
    return EMUL_BIN


def write(tmp_path, name, data):
    p = tmp_path / name
    p.write_bytes(data)
    return str(p)


def reference(path, k, lo, hi, include_all):
    opts = ["--min_threshold=%d" % lo, "--max_threshold=%d" % hi] + (["--all"] if include_all else [])
    return subprocess.run([REF_BIN] + opts + [str(k), path], check=True, capture_output=True).stdout


def emulated(path, k, lo, hi, include_all):
    return subprocess.run([emul_bin(), str(k), str(lo), str(hi), str(int(include_all)), path], check=True,
                          capture_output=True).stdout


_selected = {}


def selected(name):
    """synthetic STR-locus reads through the select stage's oracle: what the merge stage is fed"""
    if name not in _selected:
        args = dict((n, a) for n, a, _ in SYNTH)[name]
        raw = util.str_locus_reads(args[0], args[1], args[2], args[3], shared_flank=args[4])
        _selected[name] = oracle.select_buffer(raw)
    return _selected[name]


needs_ref = pytest.mark.skipif(not os.path.exists(REF_BIN), reason="oracle/_ref/downstream not built (oracle/build_downstream.sh)")


# ---------------------------------------------------------------- CPU: host sources of the product

@pytest.mark.parametrize("case", GOLDENS, ids=[g[0] for g in GOLDENS])
def test_host_build_matches_the_reference_goldens(case, tmp_path):
    name, src, k, lo, hi, include_all = case
    path = write(tmp_path, "reads.str.fq", util.gz_read(src))
    want = golden(name)
    assert emulated(path, k, lo, hi, include_all) == want
    if os.path.exists(REF_BIN):
        assert reference(path, k, lo, hi, include_all) == want           # the golden is what the reference prints here, too


def test_readme_known_answer_is_in_the_golden():
    """README.md:265-287: one polymorphic block for the shipped test_data"""
    assert golden("merge_readme_k27.txt").split(b"\n")[0] == b"@Block1\tTC\t7,5\t60\t74"


@needs_ref
@pytest.mark.parametrize("case", SYNTH, ids=[s[0] for s in SYNTH])
def test_host_build_matches_the_reference_binary(case, tmp_path):
    name, _, runs = case
    path = write(tmp_path, "reads.str.fq", selected(name))
    blocks = 0
    for k, lo, hi, include_all in runs:
        want = reference(path, k, lo, hi, include_all)
        assert emulated(path, k, lo, hi, include_all) == want, (k, lo, hi, include_all)
        blocks += want.count(b"@Block")
    assert blocks > 100


def test_host_build_input_edges(tmp_path):
    """only the end of the text ends the input: a blank line where a record should start fails the reference's assertion
    (fastq_seq.c:52), an unterminated last line its reader (utilities.c:205), half a record its record checks -- all
    are errors here too, with nothing printed"""
    sel = util.gz_read("testdata_readme.out.gz")
    cases = (("blank_line_at_the_end", sel + b"\n", "27", b"record 120 is not a 4-line FASTQ record"),
             ("records_behind_a_blank_line", sel + b"\n@x\nACGT\n+\nIIII\n", "27", b"record 120 is not a 4-line FASTQ record"),
             ("no_final_newline", sel[:-1], "27", b"last line is not terminated"),
             ("half_a_record", sel + b"@x\nACGT\n", "27", b"record 120 is not a 4-line FASTQ record"),
             ("cut_inside_a_record", b"\n".join(sel.split(b"\n")[:10]) + b"\n", "27", b"record 3 is not a 4-line FASTQ record"),
             ("not_the_select_output", util.gz_read("testdata_1.fq.gz")[:4000].rsplit(b"\n@", 1)[0] + b"\n", "27", b"error in parsing read name of record 1"),
             ("flank_shorter_than_klength", sel, "31", b"a flank is shorter than klength 31"))
    for name, data, k, what in cases:
        path = write(tmp_path, name + ".fq", data)
        r = subprocess.run([emul_bin(), k, "2", "10000", "1", path], capture_output=True)
        assert r.returncode != 0 and r.stdout == b"" and what in r.stderr, (name, r.stderr)
        if os.path.exists(REF_BIN):                              # the reference does not survive these either
            rr = subprocess.run([REF_BIN, "--all", "--min_threshold=2", k, path], capture_output=True)
            assert rr.returncode != 0 and rr.stdout == b"", name
    assert emulated(write(tmp_path, "empty.fq", b""), 27, 2, 10000, True) == b""


def test_table_order_replay_is_exercised():
    """the order of the printed blocks is the reference's hash-table order: more blocks than the 32
    starting buckets, so the replay has grown and rehashed the table several times"""
    assert golden("merge_default_k27_all_min2.txt").count(b"@Block") == 77


# ---------------------------------------------------------------- GPU: through the C ABI

@pytest.fixture(scope="module")
def capi():
    from baitstr_b200 import capi as m
    assert m.device_count() > 0, "no CUDA device: the product path cannot run (no fallback)"
    return m


@pytest.mark.gpu
@pytest.mark.parametrize("case", GOLDENS, ids=[g[0] for g in GOLDENS])
def test_gpu_matches_the_reference_goldens(capi, case):
    name, src, k, lo, hi, include_all = case
    with capi.Context() as ctx:
        out, st = ctx.merge_reads(util.gz_read(src), k, lo, hi, include_all)
    assert out == golden(name)
    assert st["printed"] == out.count(b"@Block") and st["chains"] <= st["keys"] and st["blocks"] <= st["keys"]


@pytest.mark.gpu
@pytest.mark.parametrize("case", SYNTH, ids=[s[0] for s in SYNTH])
def test_gpu_matches_the_reference_binary(capi, case, tmp_path):
    name, _, runs = case
    sel = selected(name)
    path = write(tmp_path, "reads.str.fq", sel)
    with capi.Context() as ctx:
        for k, lo, hi, include_all in runs:
            out, st = ctx.merge_reads(sel, k, lo, hi, include_all)
            want = reference(path, k, lo, hi, include_all) if os.path.exists(REF_BIN) else emulated(path, k, lo, hi, include_all)
            assert out == want, (k, lo, hi, include_all)
            assert st["reads"] == sel.count(b"\n") // 4


@pytest.mark.gpu
def test_gpu_chains_that_outgrow_the_first_geometry_are_redone(capi, tmp_path, monkeypatch):
    """the kernel first runs with shared memory sized for the input's longest flank; a chain that outgrows it is walked
    again with the largest geometry.  Forced here (BSTR_MERGE_MAXF): flanks of up to 40 / 70 bases only in the first launch"""
    sel = selected("150bp")
    path = write(tmp_path, "reads.str.fq", sel)
    want = reference(path, 27, 2, 10000, True) if os.path.exists(REF_BIN) else emulated(path, 27, 2, 10000, True)
    for maxf in ("40", "70", "383"):
        monkeypatch.setenv("BSTR_MERGE_MAXF", maxf)
        with capi.Context() as ctx:
            assert ctx.merge_reads(sel, 27, 2, 10000, True)[0] == want, maxf
            assert ctx.merge_reads(util.gz_read("testdata_default.out.gz"), 27, 2, 10000, True)[0] == golden("merge_default_k27_all_min2.txt"), maxf


@pytest.mark.gpu
def test_gpu_select_then_merge_on_the_gpu(capi, tmp_path):
    """both stages on the GPU: this repository's select output fed to its merge stage equals the
    reference's merge of the same reads"""
    raw = util.str_locus_reads(120, 50, 150, 11)
    with capi.Context() as ctx:
        sel = ctx.scan_batch(raw)[0]
        assert sel == oracle.select_buffer(raw)
        out, st = ctx.merge_reads(sel, 27, 3, 10000, True)
    path = write(tmp_path, "reads.str.fq", sel)
    want = reference(path, 27, 3, 10000, True) if os.path.exists(REF_BIN) else emulated(path, 27, 3, 10000, True)
    assert out == want and out.count(b"@Block") > 20


@pytest.mark.gpu
def test_gpu_errors_are_loud(capi):
    sel = util.gz_read("testdata_readme.out.gz")             # select ran with -f 29
    with capi.Context() as ctx:
        with pytest.raises(capi.BstrError) as ei:            # the reference reads outside the read (and crashes) here
            ctx.merge_reads(sel, 31)
        assert "flank is shorter than klength" in str(ei.value)
        with pytest.raises(capi.BstrError) as ei:            # its assertion :468 stops the reference
            ctx.merge_reads(sel, 27, 1, 10000, True)
        assert "assertion" in str(ei.value)
        with pytest.raises(capi.BstrError) as ei:            # not the select stage's output: no annotated names
            ctx.merge_reads(util.gz_read("testdata_1.fq.gz")[:4000].rsplit(b"\n@", 1)[0] + b"\n", 27)
        assert "parsing read name" in str(ei.value)
        assert ctx.merge_reads(b"", 27) == (b"", dict(reads=0, keys=0, chains=0, blocks=0, printed=0, longest_chain=0))
        assert ctx.merge_reads(sel, 27)[0] == golden("merge_readme_k27.txt")     # the context is fine afterwards


# ---------------------------------------------------------------- the command line (bin/merge_STR_reads)

CLI = os.path.join(ROOT, "bin", "merge_STR_reads")
USAGE_MD5 = "05e69bf00e36bbbf150659219c282dad"      # md5 of `merge_STR_reads --help` of the reference (0.3.1), stdout


def run_cli(*args):
    import sys
    return subprocess.run([sys.executable, CLI] + list(args), capture_output=True)


def test_cli_usage_is_the_reference_text():
    """PrintSimpleUsageString (clparsing.c:268-308) with the options of merge_STR_reads.c:524-531"""
    import hashlib
    r = run_cli("--help")
    assert r.returncode == 0 and hashlib.md5(r.stdout).hexdigest() == USAGE_MD5
    if os.path.exists(REF_BIN):
        assert subprocess.run([REF_BIN, "--help"], capture_output=True).stdout == r.stdout
    r = run_cli()                                   # too few arguments: the usage text and EXIT_FAILURE (:541-544)
    assert r.returncode == 1 and hashlib.md5(r.stdout).hexdigest() == USAGE_MD5
    r = run_cli("--bogus", "27", "x.fq")            # clparsing.c:234
    assert r.returncode == 1 and r.stdout == b"" and b"unknown option bogus" in r.stderr
    r = run_cli("--min_threshold=abc", "27", "x.fq")
    assert r.returncode == 1 and b"min_threshold: Expected unsigned int" in r.stderr


@pytest.mark.gpu
def test_cli_matches_the_reference_binary(tmp_path):
    for name, src, k, lo, hi, include_all in GOLDENS:
        path = write(tmp_path, "reads.str.fq", util.gz_read(src))
        opts = ["--min_threshold=%d" % lo, "--max_threshold=%d" % hi] + (["--all"] if include_all else ["--noall"])
        r = run_cli(*(opts[:1] + [str(k)] + opts[1:] + [path]))     # options anywhere on the line, as the reference allows
        assert r.returncode == 0 and r.stdout == golden(name), name
        nreads = util.gz_read(src).count(b"\n") // 4
        lines = r.stderr.decode().splitlines()
        assert lines[0].endswith("]\tProcessing read number 1: " + util.gz_read(src).split(b"\n", 1)[0][1:].decode())
        assert lines[-1].endswith("]\tProcessed %d reads." % nreads)
    # gzopen reads gzip and plain alike (fastq_seq.c:9); an even klength is lowered by one with a warning (:557-560)
    gz = tmp_path / "reads.str.fq.gz"
    gz.write_bytes(open(os.path.join(ROOT, "tests", "golden", "testdata_readme.out.gz"), "rb").read())
    r = run_cli("28", str(gz))
    assert r.returncode == 0 and r.stdout == golden("merge_readme_k27.txt") and b"using 27" in r.stderr
    r = run_cli("--progress=50", "27", str(gz))
    assert r.stderr.count(b"Processing read number") == 3 and b"Processing read number 101: " in r.stderr
