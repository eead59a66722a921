"""Installed layout and the downstream acceptance test of SURVEY.md section 4 / 8(c).

The reference's only end-to-end known answer (README.md:265-287, test_data/Makefile:5-12): its
select stage with `-i -n 3 -f 29` on the two shipped files, piped through its own
`merge_STR_reads 27` and `extend_STR_reads 4000 20 27`, yields exactly one polymorphic block and
one contig.  The downstream binaries are the reference's own sources compiled by
oracle/build_downstream.sh into oracle/_ref/downstream (test infrastructure).
  not gpu: the recipe and the expected lines, with the reference-generated select golden as input;
           `make install` produces the reference's installed layout under its module names
  gpu:     the same pipeline fed by THIS repo's bin/select_STR_reads
"""
import os
import subprocess
import sys

import pytest

from tests import util

ROOT = util.ROOT
DOWN = os.path.join(ROOT, "oracle", "_ref", "downstream")
MERGED = "@Block1\tTC\t7,5\t60\t74"
CONTIG = ">Block1\tTC:7,5:191:205"


def _downstream():
    exe = [os.path.join(DOWN, p) for p in ("merge_STR_reads", "extend_STR_reads")]
    if not all(os.path.exists(e) for e in exe):
        if not os.path.isdir("/root/reference/src"):
            pytest.skip("oracle/_ref/downstream not built and no reference tree here")
        subprocess.check_call(["bash", os.path.join(ROOT, "oracle", "build_downstream.sh")], stdout=subprocess.DEVNULL)
    return exe


def _pipeline(tmp_path, selected):
    merge, extend = _downstream()
    reads = []
    for i in (1, 2):
        p = tmp_path / ("Illumina_100_500_%d.fq" % i)
        p.write_bytes(util.gz_read("testdata_%d.fq.gz" % i))
        reads.append(str(p))
    sel = tmp_path / "reads.str.fq"
    sel.write_bytes(selected(reads))
    merged = subprocess.run([merge, "27", str(sel)], capture_output=True, check=True).stdout
    (tmp_path / "merged.reads.str.fq").write_bytes(merged)
    contigs = subprocess.run([extend, "4000", "20", "27", str(tmp_path / "merged.reads.str.fq")] + reads,
                             capture_output=True, check=True).stdout
    return merged.decode().splitlines(), contigs.decode().splitlines()


def _check(merged, contigs):
    assert [l for l in merged if l.startswith("@")] == [MERGED]          # one polymorphic STR location
    assert [l for l in contigs if l.startswith(">")] == [CONTIG]         # one contig


def test_reference_pipeline_known_answer(tmp_path):
    """Pins the recipe: the reference's select output (golden) through its merge and extend."""
    _check(*_pipeline(tmp_path, lambda reads: util.gz_read("testdata_readme.out.gz")))


def test_make_install_layout(tmp_path):
    """`make install PREFIX=...`: the names of /root/reference/Makefile:4-11 (minus the two C++
    stages), importable under the reference's module names, and loud without a GPU."""
    prefix = str(tmp_path / "bin")
    subprocess.check_call(["make", "-s", "-C", ROOT, "install", "PREFIX=" + prefix], stdout=subprocess.DEVNULL)
    have = set(os.listdir(prefix))
    assert {"select_STR_reads", "merge_STR_reads", "extend_STR_reads", "VERSION", "select_STR_reads.so", "fastq.so", "libbaitstr_select.so", "baitstr_b200"} <= have
    code = ("import sys; sys.path.insert(0, %r)\n"
            "from select_STR_reads import *\n"                      # /root/reference/src/select_STR_reads:61
            "import select_STR_reads as m, fastq\n"
            "assert m.__file__.startswith(%r), m.__file__\n"
            "names = ['select_str_reads', 'ProcessSequence', 'MatchPatternsAndReportBest', 'ReverseComplement', 'PrintSequence']\n"
            "assert all(callable(getattr(m, n)) for n in names)\n"
            "assert fastq.fastq and fastq.fastqsequence\n"
            "print(m.ReverseComplement('ACGTNacgtn.x'))\n"
            "from baitstr_b200 import capi\n"
            "if capi.device_count() == 0:\n"
            "    try:\n"
            "        select_str_reads(['/dev/null'], 2, 31, False, None, [2, 3, 4, 5, 6], False, False, True, 2, 100000)\n"
            "    except m.BstrError as e:\n"
            "        print('LOUD', e)\n"
            "else:\n"
            "    print('LOUD (gpu present)')\n") % (prefix, prefix)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=str(tmp_path))
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.splitlines()[0] == "x.nacgtNACGT" and "LOUD" in r.stdout
    r = subprocess.run([sys.executable, os.path.join(prefix, "select_STR_reads"), "-v"], capture_output=True, text=True)
    assert r.returncode == 0 and r.stderr.splitlines()[:2] == ["Program: select_STR_reads", "Version: 0.3.1"]


def test_fastq_module_reads_like_the_reference(tmp_path):
    """bin/fastq.so: records, upper-casing, '.gz' by suffix, the asserts of src/fastq.pyx:72-94."""
    import gzip
    prefix = os.path.join(ROOT, "bin")
    if not os.path.exists(os.path.join(prefix, "fastq.so")):
        subprocess.check_call(["make", "-s", "-C", ROOT, "install"], stdout=subprocess.DEVNULL)
    raw = b"@r1 x\tc\nacgtn.ACGT\n+r1\nIIIIIIIIII\n@r2\nTT\n+\nII"
    (tmp_path / "a.fq").write_bytes(raw)
    (tmp_path / "a.fq.gz").write_bytes(gzip.compress(raw))
    (tmp_path / "bad.fq").write_bytes(b"@r\nACRT\n+\nIIII\n")
    code = ("import sys; sys.path.insert(0, %r); import fastq\n"
            "for fn in ('a.fq', 'a.fq.gz'):\n"
            "    recs = [(r.fastqsequence.name, r.fastqsequence.seq, r.fastqsequence.qual) for r in fastq.fastq(fn)]\n"
            "    assert recs == [('r1 x\\tc', 'ACGTN.ACGT', 'IIIIIIIIII'), ('r2', 'TT', 'I')], recs\n"
            "try:\n"
            "    list(fastq.fastq('bad.fq'))\n"
            "except AssertionError as e:\n"
            "    print('ASSERT', e)\n") % prefix
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=str(tmp_path))
    assert r.returncode == 0 and "ASSERT read r should only have ACGTN" in r.stdout, r.stderr[-2000:] + r.stdout


@pytest.mark.gpu
def test_gpu_select_feeds_reference_merge_and_extend(tmp_path):
    """This repo's CLI (README flags of test_data/Makefile:6) -> the reference's merge -> extend."""
    def selected(reads):
        r = subprocess.run([sys.executable, os.path.join(ROOT, "bin", "select_STR_reads"), "-i", "-n", "3", "-f", "29"] + reads,
                           capture_output=True)
        assert r.returncode == 0, r.stderr[-2000:]
        assert r.stdout == util.gz_read("testdata_readme.out.gz")       # and byte-identical to the reference's select
        return r.stdout
    _check(*_pipeline(tmp_path, selected))


@pytest.mark.gpu
def test_gpu_select_and_merge_feed_reference_extend(tmp_path):
    """README.md:265-287 with BOTH first stages from this repository: bin/select_STR_reads | bin/merge_STR_reads 27,
    then the reference's extend_STR_reads -- the same block and the same contig."""
    _, extend = _downstream()
    reads = []
    for i in (1, 2):
        p = tmp_path / ("Illumina_100_500_%d.fq" % i)
        p.write_bytes(util.gz_read("testdata_%d.fq.gz" % i))
        reads.append(str(p))
    sel = subprocess.run([sys.executable, os.path.join(ROOT, "bin", "select_STR_reads"), "-i", "-n", "3", "-f", "29"] + reads,
                         capture_output=True, check=True).stdout
    (tmp_path / "reads.str.fq").write_bytes(sel)
    merged = subprocess.run([sys.executable, os.path.join(ROOT, "bin", "merge_STR_reads"), "27", str(tmp_path / "reads.str.fq")],
                            capture_output=True, check=True).stdout
    (tmp_path / "merged.reads.str.fq").write_bytes(merged)
    contigs = subprocess.run([extend, "4000", "20", "27", str(tmp_path / "merged.reads.str.fq")] + reads, capture_output=True,
                             check=True).stdout
    _check(merged.decode().splitlines(), contigs.decode().splitlines())


@pytest.mark.gpu
def test_gpu_whole_pipeline_from_this_repository(tmp_path):
    """README.md:265-287 with all three stages from this repository (test_data/Makefile:5-12):
    bin/select_STR_reads -i -n 3 -f 29 | bin/merge_STR_reads 27 | bin/extend_STR_reads 4000 20 27 --
    the block and the contig of the reference's walk-through, byte for byte its goldens."""
    reads = []
    for i in (1, 2):
        p = tmp_path / ("Illumina_100_500_%d.fq" % i)
        p.write_bytes(util.gz_read("testdata_%d.fq.gz" % i))
        reads.append(str(p))
    run = lambda exe, *args: subprocess.run([sys.executable, os.path.join(ROOT, "bin", exe)] + list(args), capture_output=True, check=True).stdout
    (tmp_path / "reads.str.fq").write_bytes(run("select_STR_reads", "-i", "-n", "3", "-f", "29", *reads))
    merged = run("merge_STR_reads", "27", str(tmp_path / "reads.str.fq"))
    (tmp_path / "merged.reads.str.fq").write_bytes(merged)
    contigs = run("extend_STR_reads", "4000", "20", "27", str(tmp_path / "merged.reads.str.fq"), *reads)
    _check(merged.decode().splitlines(), contigs.decode().splitlines())
    golden = lambda n: open(os.path.join(ROOT, "tests", "golden", n), "rb").read()
    assert merged == golden("merge_readme_k27.txt") and contigs == golden("extend_readme.txt")
