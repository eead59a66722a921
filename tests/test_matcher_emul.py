"""CPU fuzz of the DEVICE matcher source (matcher.cuh built for the host, tests/emul) against
the oracle.  Catches logic errors in the bit-parallel formulation without a GPU; the GPU
parity tests proper are in tests/test_gpu_parity.py."""
import random

import numpy as np
import pytest

from oracle import oracle
from tests import util
from tests.emul import emul


def expect(seq, **kw):
    """(fstart, rstart, fck, rck) the k_match_reads convention gives for an upper-cased read."""
    s = seq.upper()
    fwd = oracle.match_strand(s, **kw)
    if fwd is None:
        return (0, 0, 0, 0)
    rc = oracle.match_strand(oracle.revcomp(s), **kw)
    fck = fwd[1] << 3 | len(fwd[0])
    if rc is None:
        return (fwd[2], 0, fck, 0)
    return (fwd[2], rc[2], fck, rc[1] << 3 | len(rc[0]))


def rand_read(rng, L, alpha):
    s = [rng.choice(alpha) for _ in range(L)]
    for _ in range(rng.choice([0, 1, 1, 2, 3])):
        k = rng.randint(1, 6)
        m = [rng.choice(alpha) for _ in range(k)]
        c = rng.randint(2, 14)
        p = rng.randint(0, max(0, L - 1))
        for j in range(k * c):
            if p + j < L:
                s[p + j] = m[j % k]
    if rng.random() < 0.2 and L:
        s[rng.randrange(L)] = rng.choice("N.")
    s = "".join(s)
    return s.lower() if rng.random() < 0.1 else s


def check(seqs, force_w=0, **kw):
    got = emul.match_reads([s.encode("latin-1") for s in seqs], force_w=force_w, **kw)
    for s, g in zip(seqs, got):
        want = expect(s, **kw)
        assert tuple(int(x) for x in g) == want, (s, kw, force_w, tuple(int(x) for x in g), want)


def test_known_answers():
    for seq, kw, fwd, rc in util.KNOWN:
        args = dict(n=kw.get("n", 2), f=kw.get("f", 0), ks=kw.get("ks"),
                    rmhm=kw.get("rmhm", False), altalgo=kw.get("altalgo", True))
        for w in (0, 2, 4, 6, 8, 9):
            check([seq], force_w=w, **args)


@pytest.mark.parametrize("seed", range(6))
def test_fuzz_vs_oracle(seed):
    rng = random.Random(1000 + seed)
    for it in range(60):
        n = rng.choice([1, 2, 2, 3, 4, 5])
        f = rng.choice([0, 1, 5, 6, 7, 10, 31, 41])
        ks = rng.choice([None, [2], [3], [6], [2, 3], [4, 5], [2, 6], [3, 4, 5, 6]])
        rmhm, alt = rng.random() < 0.4, rng.random() < 0.75
        alpha = rng.choice(["ACGT", "AC", "ACGTN", "ACGT.", "A", "AT"])
        seqs = []
        for _ in range(80):
            L = rng.choice([0, 1, 3, 12, 31, 32, 33, 63, 64, 65, 76, 100, 101, 127, 128, 129, 150,
                            200, 250, 255, 256, 257, 300, 400])
            seqs.append(rand_read(rng, L, alpha))
        check(seqs, n=n, f=f, ks=ks, rmhm=rmhm, altalgo=alt)
        # the 6-word window (129..192 window bases) through the WMAX = 6 instantiation
        check([s for s in seqs if len(s) <= 192 + 2 * max(0, (f if alt else 0) - 6)][:40], force_w=6,
              n=n, f=f, ks=ks, rmhm=rmhm, altalgo=alt)
        if it % 4 == 0:
            check([s for s in seqs if len(s) <= 64 + 2 * max(0, f - 6)][:20], force_w=rng.choice([4, 6, 8, 9]),
                  n=n, f=f, ks=ks, rmhm=rmhm, altalgo=alt)


def test_fuzz_goldens_from_reference():
    cases = util.fuzz_cases()
    by = {}
    for c in cases:
        by.setdefault((c["n"], c["f"], tuple(c["ks"]), c["rmhm"], c["alt"]), []).append(c)
    for (n, f, ks, rmhm, alt), cs in by.items():
        got = emul.match_reads([c["s"].encode("latin-1") for c in cs], n=n, f=f, ks=list(ks),
                               rmhm=rmhm, altalgo=alt)
        for c, g in zip(cs, got):
            fwd, rc = c["fwd"], c["rc"]
            if fwd is None:
                want = (0, 0, 0, 0)
            elif rc is None:
                want = (fwd[2], 0, fwd[1] << 3 | len(fwd[0]), 0)
            else:
                want = (fwd[2], rc[2], fwd[1] << 3 | len(fwd[0]), rc[1] << 3 | len(rc[0]))
            assert tuple(int(x) for x in g) == want, c


def test_swar_base_validator_exhaustive():
    """fastq.pyx:62-67: every byte value, every alignment and position in a word."""
    ok = set(b"acgtnACGTN.")
    for mis in range(4):
        for L in (1, 2, 3, 4, 5, 7, 8, 9, 33):
            assert not emul.bases_invalid(b"ACGTNacgtn."[:L] * 1 if L <= 11 else b"ACGT.nN" * 5, mis)
            for pos in {0, L // 2, L - 1}:
                for b in range(256):
                    s = bytearray((b"ACGTN.acgtn" * 4)[:L])
                    s[pos] = b
                    assert emul.bases_invalid(bytes(s), mis) == (b not in ok), (mis, L, pos, b)
    assert not emul.bases_invalid(b"", 1)
