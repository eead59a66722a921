"""CPU tests: the C restatement (oracle/str_oracle.c) against the reference's goldens.

The goldens under tests/golden/ were produced by the UNMODIFIED reference
(tests/golden/make_golden.py); when oracle/_ref is present the live reference is fuzzed too.
"""
import hashlib
import os
import sys

import pytest

from oracle import oracle
from tests import util


def test_known_answers():
    for seq, kw, fwd, rc in util.KNOWN:
        args = dict(n=kw.get("n", 2), f=kw.get("f", 0), ks=kw.get("ks"),
                    rmhm=kw.get("rmhm", False), altalgo=kw.get("altalgo", True))
        assert oracle.match_strand(seq, **args) == fwd, (seq, kw)
        assert oracle.match_strand(oracle.revcomp(seq), **args) == rc, (seq, kw)


@pytest.mark.parametrize("tag", ["default", "readme", "n3r", "k23f41", "x"])
def test_testdata_goldens(tag):
    man = util.manifest()["testdata_" + tag]
    kw = util.flags_to_kwargs(man["flags"])
    out = b"".join(oracle.select_buffer(util.gz_read("testdata_%d.fq.gz" % i), **kw) for i in (1, 2))
    assert hashlib.md5(out).hexdigest() == man["md5"]
    assert out == util.gz_read("testdata_%s.out.gz" % tag)


def test_survey_md5s():
    man = util.manifest()
    assert man["testdata_default"]["md5"] == "79a4e5db8956c8e61f2d86bca6fbea88"
    assert man["testdata_readme"]["md5"] == "51248b2b3f490241db2468c60a933057"


def test_fuzz_goldens():
    cases = util.fuzz_cases()
    assert len(cases) >= 6000
    for c in cases:
        args = dict(n=c["n"], f=c["f"], ks=c["ks"], rmhm=c["rmhm"], altalgo=c["alt"])
        fwd = oracle.match_strand(c["s"], **args)
        rc = oracle.match_strand(oracle.revcomp(c["s"]), **args)
        assert fwd == (tuple(c["fwd"]) if c["fwd"] else None), c
        assert rc == (tuple(c["rc"]) if c["rc"] else None), c


def test_edge_cases():
    for name, case in util.edge_cases().items():
        raw = case["input"].encode("latin-1")
        if name == "high_bytes_qual":
            continue   # the Py3 shim build of the reference cannot carry bytes > 127 (char* ascii)
        if case["ok"]:
            assert oracle.select_buffer(raw, n=2, f=3) == case["stdout"].encode("latin-1"), name
        else:
            with pytest.raises(oracle.OracleError):
                oracle.select_buffer(raw, n=2, f=3)


def test_live_reference_fuzz():
    """Fuzz against the live reference build when it is present (build container / GPU box)."""
    try:
        from oracle.ref_select import load_reference
        ref = load_reference()
    except Exception:
        pytest.skip("oracle/_ref not built")
    import random
    rng = random.Random(7)
    for it in range(3000):
        L = rng.choice([20, 40, 64, 100, 150])
        alpha = rng.choice(["ACGT", "AC", "ACGTN", "ACGT."])
        s = "".join(rng.choice(alpha) for _ in range(L))
        if rng.random() < 0.5:
            k = rng.randint(1, 6); m = "".join(rng.choice(alpha) for _ in range(k))
            p = rng.randint(0, L - 1); rep = (m * 20)[:rng.randint(2 * k, 12 * k)]
            s = (s[:p] + rep + s[p + len(rep):])[:L]
        n, f = rng.choice([2, 3, 4]), rng.choice([0, 5, 10, 31])
        ks = rng.choice([[2, 3, 4, 5, 6], [2, 3], [5], [4, 6]])
        rmhm, alt = rng.random() < 0.3, rng.random() < 0.7
        pats = [r'(?=((?=(.{%d}))\2{%d,}))' % (k, n) for k in ks]
        for strand in (s, ref.ReverseComplement(s)):
            m = ref.MatchPatternsAndReportBest(strand, pats, rmhm, f, False, alt)
            want = None if m is None else (m[3], int(m[1]), m[4], m[5])
            assert oracle.match_strand(strand, n=n, f=f, ks=ks, rmhm=rmhm, altalgo=alt) == want
