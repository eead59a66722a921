#!/usr/bin/env python3
"""Regenerate tests/golden/* from the UNMODIFIED reference (oracle/_ref).

Run in the build container (needs /root/reference for the shipped test_data inputs):
    python oracle/build_ref.py && python tests/golden/make_golden.py
Writes
  testdata_{1,2}.fq.gz              the reference's only fixture (test_data/Illumina_100_500_{1,2}.fq)
  testdata_default.out.gz           reference output, default flags           (BASELINE config C1)
  testdata_readme.out.gz            reference output, -i -n 3 -f 29           (test_data/Makefile:6)
  fuzz_cases.json.gz                strand-matcher answers of the reference on seeded random strings
  edge_cases.json                   whole-stage outputs of the reference on hand-written FASTQ snippets
  MANIFEST.json                     md5 of every golden output
"""
import gzip
import hashlib
import json
import os
import random
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
TD = "/root/reference/test_data"
RUNNER = os.path.join(ROOT, "oracle", "ref_select.py")


def run_ref(flags, files):
    r = subprocess.run([sys.executable, RUNNER] + flags + files, stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE)
    return r.returncode, r.stdout


def gzwrite(path, data):
    with open(path, "wb") as raw:
        with gzip.GzipFile(fileobj=raw, mode="wb", mtime=0, filename="") as fh:
            fh.write(data)


def fuzz_cases(ref):
    return fuzz_block(ref, 0xBA175712, 6000, [2, 2, 2, 3, 3, 5], [0, 0, 3, 5, 8, 31, 31, 41]) + \
        fuzz_block(ref, 0xBA175713, 3000, [0, 1, 1, 2, 3, 5], [0, 1, 5, 6, 7, 8, 31, 41])
    # second block (round 2): the n <= 1 branch (every position is a one-copy match) and the
    # window-origin boundary f = 6 / 7 (g0 = f - 6 in csrc/matcher.cuh) pinned to the reference too


def fuzz_block(ref, seed, count, n_choices, f_choices):
    rng = random.Random(seed)
    cases = []
    alphabets = ["ACGT", "ACGT", "AC", "ACGTN", "ACGT.", "ACGTN."]
    for i in range(count):
        alpha = alphabets[i % len(alphabets)]
        L = rng.choice([0, 1, 5, 13, 31, 32, 33, 63, 64, 65, 100, 100, 100, 150, 250])
        s = [rng.choice(alpha) for _ in range(L)]
        for _ in range(rng.randint(0, 3)):       # plant repeats, some straddling the flanks
            k = rng.randint(1, 6)
            motif = [rng.choice(alpha) for _ in range(k)]
            c = rng.randint(2, 14)
            pos = rng.randint(0, max(0, L - 1))
            for j in range(k * c + rng.randint(0, k)):
                if pos + j < L:
                    s[pos + j] = motif[j % k]
        s = "".join(s)
        n = rng.choice(n_choices)
        f = rng.choice(f_choices)
        ks = rng.choice([[2, 3, 4, 5, 6]] * 3 + [[2], [3], [2, 3], [4], [5, 6], [6], [2, 4, 6]])
        rmhm = rng.random() < 0.3
        alt = rng.random() < 0.75
        patterns = [r'(?=((?=(.{%d}))\2{%d,}))' % (k, n) for k in ks]
        ans = []
        for strand in (s, ref.ReverseComplement(s)):
            m = ref.MatchPatternsAndReportBest(strand, patterns, rmhm, f, False, alt)
            ans.append(None if m is None else [m[3], int(m[1]), m[4], m[5]])
        cases.append({"s": s, "n": n, "f": f, "ks": ks, "rmhm": rmhm, "alt": alt,
                      "fwd": ans[0], "rc": ans[1]})
    return cases


EDGE_INPUTS = {
    "comment_header": "@r1 x\tcomment\nTTACACACACGTT\n+r1\nIIIIIIIIIIIII\n",
    "lowercase_dot": "@r2\nttacacacac.tg\n+\nIIIIIIIIIIIII\n",
    "no_final_newline": "@r3\nTTACACACACGTT\n+\nIIIIIIIIIIIIJ",
    "qual_short": "@r4\nTTACACACACGTT\n+\nIII\n",
    "qual_long": "@r5\nTTACACACACGTT\n+\nIIIIIIIIIIIIIIIIIIIIIIII\n",
    "empty_seq": "@r6\n\n+\n\n@r7\nTTACACACACGTT\n+\nIIIIIIIIIIIII\n",
    "empty_file": "",
    "two_records_N": "@a\nTTACACNCACGTT\n+\nIIIIIIIIIIIII\n@b\nGGCATATATATCG\n+\nJJJJJJJJJJJJJ\n",
    "trailing_blank_line": "@r8\nTTACACACACGTT\n+\nIIIIIIIIIIIII\n\n",
    "crlf": "@r9\r\nTTACACACACGTT\r\n+\r\nIIIIIIIIIIIII\r\n",
    "bad_base": "@r10\nTTACACRCACGTT\n+\nIIIIIIIIIIIII\n",
    "truncated": "@r11\nTTACACACACGTT\n",
    "no_at": "r12\nTTACACACACGTT\n+\nIIIIIIIIIIIII\n",
    "no_plus": "@r13\nTTACACACACGTT\n-\nIIIIIIIIIIIII\n",
    "header_only_at": "@\nTTACACACACGTT\n+\nIIIIIIIIIIIII\n",
    "high_bytes_qual": "@r14\nTTACACACACGTT\n+\n\xe9\xff~}|{zyxwvut\n",
}


def edge_cases():
    out = {}
    for name, text in EDGE_INPUTS.items():
        with tempfile.NamedTemporaryFile("wb", suffix=".fq", delete=False) as fh:
            fh.write(text.encode("latin-1"))
            path = fh.name
        rc, stdout = run_ref(["-n", "2", "-f", "3"], [path])
        os.unlink(path)
        # Python 3 prints latin-1 text as UTF-8; undo that so goldens are raw bytes.
        try:
            stdout = stdout.decode("utf-8").encode("latin-1")
        except (UnicodeDecodeError, UnicodeEncodeError):
            pass
        out[name] = {"input": text, "ok": rc == 0,
                     "stdout": stdout.decode("latin-1") if rc == 0 else None}
    return out


def main():
    from oracle.ref_select import load_reference
    ref = load_reference()
    manifest = {}
    files = []
    for i in (1, 2):
        data = open(os.path.join(TD, "Illumina_100_500_%d.fq" % i), "rb").read()
        gzwrite(os.path.join(HERE, "testdata_%d.fq.gz" % i), data)
        files.append(os.path.join(TD, "Illumina_100_500_%d.fq" % i))
    for tag, flags in (("default", []), ("readme", ["-i", "-n", "3", "-f", "29"]),
                       ("n3r", ["-n", "3", "-r", "-f", "20"]), ("k23f41", ["--2mer", "--3mer", "-f", "41"]),
                       ("x", ["-x", "-f", "10"])):
        rc, out = run_ref(flags, files)
        assert rc == 0
        gzwrite(os.path.join(HERE, "testdata_%s.out.gz" % tag), out)
        manifest["testdata_%s" % tag] = {"flags": flags, "md5": hashlib.md5(out).hexdigest(),
                                         "lines": out.count(b"\n")}
    gzwrite(os.path.join(HERE, "fuzz_cases.json.gz"),
            json.dumps(fuzz_cases(ref), separators=(",", ":")).encode())
    with open(os.path.join(HERE, "edge_cases.json"), "w") as fh:
        json.dump(edge_cases(), fh, indent=1)
    with open(os.path.join(HERE, "MANIFEST.json"), "w") as fh:
        json.dump(manifest, fh, indent=1)
    print(json.dumps(manifest, indent=1))


if __name__ == "__main__":
    main()
