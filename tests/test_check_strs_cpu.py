"""`-c/--check` host logic (baitstr_b200/check_strs.py) against the stderr of the live
reference (tests/golden/check_strs.json, made by tests/golden/make_check_golden.py).

No GPU here: the annotated FASTQ the report reads is produced by the oracle (test
infrastructure); the GPU path end to end is tests/test_gpu_parity.py::test_cli_check_strs."""
import io
import json
import os

import pytest

from baitstr_b200 import check_strs
from oracle import oracle
from tests import util

CASES = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "check_strs.json")))


@pytest.mark.parametrize("ci", range(len(CASES)))
def test_report_matches_reference_stderr(ci, tmp_path):
    case = CASES[ci]
    chk = tmp_path / "check.txt"
    chk.write_text(case["check_file"])
    to_check = check_strs.read_str_list(str(chk))
    kw = util.flags_to_kwargs(case["flags"])
    err = io.StringIO()
    for i in (1, 2):
        annotated = oracle.select_buffer(util.gz_read("testdata_%d.fq.gz" % i), **kw)
        check_strs.register_annotated_fastq(annotated, to_check, err)
        err.write("Done processing testdata_%d.fq\n" % i)
    check_strs.print_performance_log(to_check, err)
    assert err.getvalue() == case["stderr"]
    # the counts the reference loses in its pool workers are kept here
    found = sum(e.found for es in to_check.values() for e in es)
    nomatch = case["stderr"].count("No match for")
    nsel = sum(oracle.select_buffer(util.gz_read("testdata_%d.fq.gz" % i), **kw).count(b"\n") // 4 for i in (1, 2))
    assert found + nomatch == nsel and found > 0


def test_read_str_list_and_identity():
    assert check_strs.read_str_list(None) == {}
    e = check_strs.Entry("ACGTACGTAC", "TTTTT")
    assert check_strs.percent_identity(e, "acgtacgtac", "TTTTT") == 100.0          # case-insensitive
    assert check_strs.percent_identity(e, "GTAC", "TTTAA") == pytest.approx(100.0 * 7 / 9)   # shorter flank decides
    with pytest.raises(ZeroDivisionError):                                         # pyx:207
        check_strs.percent_identity(e, "", "")


def test_header_names_with_tabs():
    """The read name may itself contain tabs (fastq.pyx:78 keeps the whole header line)."""
    rec = b"@r1 x\tcomment\tCA\t3\t3\t9\tTG\t3\t4\t10\nTTACACACAC.TG\n+\nIIIIIIIIIIIII\n"
    err = io.StringIO()
    check_strs.register_annotated_fastq(rec, {("CA", 3): [check_strs.Entry("TTA", "C.TG")]}, err)
    assert err.getvalue() == ""                                                    # found, nothing printed
    check_strs.register_annotated_fastq(rec, {}, err)
    assert err.getvalue().startswith("No match for TG:3:")
