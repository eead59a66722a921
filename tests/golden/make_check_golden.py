#!/usr/bin/env python3
"""Regenerate tests/golden/check_strs.json from the UNMODIFIED reference (oracle/_ref):
the stderr of `select_STR_reads -c LIST` (src/select_STR_reads.pyx:17-49, 191-292) on the
shipped test data.  Run in the build container:  python tests/golden/make_check_golden.py

The check list is derived from the reference's own output so that every branch of
RegisterSTR occurs: listed STRs with their exact flanks (found on the forward key), the same
written for the opposite strand (found on the reverse-complement key), flanks with a few
mismatches (identity below and above 90 %), several entries under one key, and entries no
read supports.  Two flag sets: the defaults, and -f 5 (STRs that start before base 20, where
the reference's left flank is sliced with a negative lower bound).
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
COMP = str.maketrans("atcgnATCGN", "tagcnTAGCN")


def run_ref(flags, files):
    r = subprocess.run([sys.executable, RUNNER, "-p", "1"] + flags + files, stdout=subprocess.PIPE,
                       stderr=subprocess.PIPE)
    assert r.returncode == 0, r.stderr[-2000:]
    return r.stdout.decode("latin-1"), r.stderr.decode("latin-1")


def mutate(rng, s, k):
    s = list(s)
    for i in rng.sample(range(len(s)), min(k, len(s))):
        s[i] = rng.choice([c for c in "ACGT" if c != s[i]])
    return "".join(s)


def check_list(rng, stdout, n_each):
    """(motif, copies, lflank, rflank) lines from selected reads of the reference's output."""
    lines = stdout.split("\n")
    recs = []
    for i in range(0, len(lines) - 3, 4):
        f = lines[i].rsplit("\t", 8)
        seq = lines[i + 1]
        recs.append((f[1], int(f[2]), int(f[3]), int(f[4]), seq))
    rng.shuffle(recs)
    out = []
    for j, (motif, c, s, e, seq) in enumerate(recs[:5 * n_each]):
        lf, rf = seq[max(0, s - 20):s], seq[e:e + 20]
        kind = j % 5
        if kind == 0:
            out.append((motif, c, lf, rf))                                   # exact, forward key
        elif kind == 1:                                                      # written for the other strand
            out.append((motif.translate(COMP)[::-1], c, rf.translate(COMP)[::-1], lf.translate(COMP)[::-1]))
        elif kind == 2:
            out.append((motif, c, mutate(rng, lf, 1), mutate(rng, rf, 2)))   # 92.5 % on full flanks
        elif kind == 3:
            out.append((motif, c, mutate(rng, lf, 6), mutate(rng, rf, 7)))   # well below 90 %
        else:
            out.append((motif, c, lf[-7:].lower(), rf[:9]))                  # short, lower-case flanks
    out.append(("ACGTAC", 9, "AAAAAAAAAAAAAAAAAAAA", "CCCCCCCCCCCCCCCCCCCC"))  # supported by nothing
    out.append((out[0][0], out[0][1], "TTTTTTTTTTTTTTTTTTTT", "GGGGGGGGGGGGGGGGGGGG"))  # second entry of a key
    return out


def main():
    rng = random.Random(0xC4EC)
    cases = []
    with tempfile.TemporaryDirectory() as tmp:
        files = []
        for i in (1, 2):
            p = os.path.join(tmp, "testdata_%d.fq" % i)
            with gzip.open(os.path.join(HERE, "testdata_%d.fq.gz" % i), "rb") as fh, open(p, "wb") as out:
                out.write(fh.read())
            files.append(p)
        for flags in ([], ["-f", "5"], ["-n", "3", "-r", "-f", "20"]):
            stdout, _ = run_ref(flags, files)
            chk = check_list(rng, stdout, 12)
            cpath = os.path.join(tmp, "check.txt")
            with open(cpath, "w") as fh:
                for m, c, lf, rf in chk:
                    fh.write("%s %d %s %s\n" % (m, c, lf, rf))
            stdout2, stderr = run_ref(flags + ["-c", cpath], files)
            assert stdout2 == stdout                                         # -c never changes stdout
            cases.append({"flags": flags, "check_file": open(cpath).read(),
                          "stderr": stderr.replace(tmp + "/", "")})
    with open(os.path.join(HERE, "check_strs.json"), "w") as fh:
        json.dump(cases, fh, indent=1)
    for c in cases:
        err = c["stderr"].split("\n")
        print(c["flags"], "stderr lines:", len(err), "no-match:", sum(l.startswith("No match") for l in err),
              "found:", sum(l.startswith("Found") for l in err))


if __name__ == "__main__":
    main()
