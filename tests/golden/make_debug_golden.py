#!/usr/bin/env python3
"""Regenerate tests/golden/debug_report.json.gz from the UNMODIFIED reference (oracle/_ref): the
stderr of `select_STR_reads -d` (src/select_STR_reads.pyx:76-87, 138-140, 157-163, 296-311), one
worker (`-p 1`: with more the reference's blocks interleave).  Run in the build container:
    python tests/golden/make_debug_golden.py
Inputs: the first 60 records of the shipped test data plus hand-made reads (N inside and
beside repeats, '.', lower case, homopolymers, repeats in the flanks, very short reads); six flag
sets including -x and -c together with -d.
"""
import gzip
import json
import os
import random
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
RUNNER = os.path.join(ROOT, "oracle", "ref_select.py")

FLAGSETS = [[], ["-x", "-f", "10"], ["-n", "3", "-r", "-f", "20"], ["-2", "-3", "-f", "41"],
            ["-x", "-r", "-f", "5", "-n", "3"], ["-n", "1", "-f", "0", "-6"], ["-f", "8", "-n", "2", "-4", "-5"]]


def hand_made(rng):
    reads = ["GGCATATATATCG", "CGTAAAAAAATGC", "CGTANANANATGC", "CGTACACNCACACTG", "CGT....TGC",
             "TTACGTAGACGTAGACGTAGACGTT", "acacacacacacacttgtgtgggcaatcagc", "A", "AC", "",
             "NNNNNNNNNNNN", "ACGTNACGTNACGTNACGTNACGTN", "TTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTTT"]
    for _ in range(40):
        L = rng.choice([20, 37, 64, 100, 130])
        s = [rng.choice("ACGT") for _ in range(L)]
        for _ in range(rng.randint(0, 3)):
            k = rng.randint(1, 6)
            m = [rng.choice("ACGTN"[: rng.choice([4, 4, 5])]) for _ in range(k)]
            c = rng.randint(2, 9)
            p = rng.randint(0, max(0, L - 1))
            for j in range(min(k * c, L - p)):
                s[p + j] = m[j % k]
        seq = "".join(s)
        reads.append(seq.lower() if rng.random() < 0.1 else seq)
    return reads


def main():
    rng = random.Random(0xDEB6)
    with gzip.open(os.path.join(HERE, "testdata_1.fq.gz"), "rb") as fh:
        shipped = b"\n".join(fh.read().split(b"\n")[:240]) + b"\n"
    reads = hand_made(rng)
    made = "".join("@hand%d extra\tfield\n%s\n+\n%s\n" % (i, s, "I" * len(s)) for i, s in enumerate(reads))
    fastq = shipped + made.encode("latin-1")
    cases = []
    with tempfile.TemporaryDirectory() as tmp:
        path = os.path.join(tmp, "debug.fq")
        open(path, "wb").write(fastq)
        chk = os.path.join(tmp, "check.txt")
        open(chk, "w").write("AT 4 GGC CG\nGAT 2 GAGAGAGACAATCCGGAC TCCGTTACTTGAGTCGCTTT\n")
        for flags in FLAGSETS + [["-c", chk, "-f", "3"]]:
            r = subprocess.run([sys.executable, RUNNER, "-p", "1", "-d"] + flags + [path],
                               stdout=subprocess.PIPE, stderr=subprocess.PIPE)
            assert r.returncode == 0, r.stderr[-2000:]
            cases.append({"flags": [f if f != chk else "CHECKFILE" for f in flags],
                          "stderr": r.stderr.decode("latin-1").replace(tmp + "/", "")})
        out = {"fastq": fastq.decode("latin-1"), "check_file": open(chk).read(), "cases": cases}
    with gzip.GzipFile(os.path.join(HERE, "debug_report.json.gz"), "wb", mtime=0) as fh:
        fh.write(json.dumps(out).encode())
    for c in cases:
        print(c["flags"], "stderr lines:", c["stderr"].count("\n"))


if __name__ == "__main__":
    main()
