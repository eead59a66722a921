"""`-d/--debug` report against the stderr of the live reference
(tests/golden/debug_report.json.gz, made by tests/golden/make_debug_golden.py).

No GPU here: the device code that enumerates the rejected regex hits (csrc/matcher.cuh:
debug_read, the body of k_debug_reads) runs on the CPU through tests/emul, and
baitstr_b200/debug_report.py renders it.  The GPU path end to end is
tests/test_gpu_parity.py::test_cli_debug_report."""
import gzip
import io
import json
import os

import pytest

from baitstr_b200 import check_strs, debug_report
from oracle import oracle
from tests import util
from tests.emul import emul

GOLD = json.loads(gzip.open(os.path.join(os.path.dirname(__file__), "golden", "debug_report.json.gz")).read())


def records(fastq):
    lines = fastq.split("\n")
    return [(lines[i][1:], lines[i + 1].upper()) for i in range(0, len(lines) - 3, 4)]


@pytest.mark.parametrize("ci", range(len(GOLD["cases"])))
def test_blocks_match_reference_stderr(ci, tmp_path):
    case = GOLD["cases"][ci]
    flags = [f for f in case["flags"] if f not in ("-c", "CHECKFILE")]
    kw = util.flags_to_kwargs(flags)
    to_check = None
    check_lines = None
    if "-c" in case["flags"]:
        chk = tmp_path / "check.txt"
        chk.write_text(GOLD["check_file"])
        to_check = check_strs.read_str_list(str(chk))
        annotated = oracle.select_buffer(GOLD["fastq"].encode("latin-1"), **kw)
        check_lines = iter(check_strs.annotated_fastq_lines(annotated, to_check))
    err = io.StringIO()
    for name, seq in records(GOLD["fastq"]):
        info, fwd, rc = emul.debug_read(seq.encode("latin-1"), n=kw["n"], f=kw["f"], ks=kw["ks"],
                                        rmhm=kw["rmhm"], altalgo=kw["altalgo"])
        err.write(debug_report.render_read(name, info, fwd, rc))
        if check_lines is not None and not (info[1] or info[5]):
            err.write(next(check_lines))
    err.write("Done processing debug.fq\n")
    if to_check is not None:
        check_strs.print_performance_log(to_check, err)
    assert err.getvalue() == case["stderr"]
