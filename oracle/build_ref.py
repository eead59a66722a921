#!/usr/bin/env python3
"""Build the UNMODIFIED reference select stage into oracle/_ref/ (test infrastructure).

The two Cython sources are compiled where they lie under /root/reference/src
(fastq.pyx, select_STR_reads.pyx); nothing is copied into the repo.  Only generated
C and the two extension modules land in oracle/_ref/ (git-ignored, but shipped to the
GPU box by gpurun).  Recipe = SURVEY.md section 8(c): language_level=2 plus
c_string_type=str / c_string_encoding=ascii so the `char*` constructor arguments of
fastq.pyx:16,49 accept Python-3 str.

Usage: python oracle/build_ref.py [--reference /root/reference]
Exit 0 and print "skipped" if the reference tree is absent (GPU box: prebuilt files
are used as they are).
"""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
MODULES = ("fastq", "select_STR_reads")


def built_paths():
    suffix = sysconfig.get_config_var("EXT_SUFFIX")
    return [os.path.join(OUT, m + suffix) for m in MODULES]


def build(reference="/root/reference", force=False):
    src = os.path.join(reference, "src")
    if not os.path.isdir(src):
        print("oracle/_ref: reference tree absent, build skipped")
        return all(os.path.exists(p) for p in built_paths())
    os.makedirs(os.path.join(OUT, "build"), exist_ok=True)
    inc = sysconfig.get_paths()["include"]
    for mod, so in zip(MODULES, built_paths()):
        pyx = os.path.join(src, mod + ".pyx")
        c = os.path.join(OUT, "build", mod + ".c")
        if not force and os.path.exists(so) and os.path.getmtime(so) >= os.path.getmtime(pyx):
            continue
        subprocess.check_call([sys.executable, "-m", "cython", "-2",
                               "-X", "c_string_type=str", "-X", "c_string_encoding=ascii",
                               "-o", c, pyx])
        subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-w", "-fwrapv",
                               "-I", inc, c, "-o", so])
    with open(os.path.join(src, "VERSION")) as fh:
        version = fh.read().strip()
    with open(os.path.join(OUT, "VERSION"), "w") as fh:   # 6 bytes of metadata, not source
        fh.write(version + "\n")
    return True


if __name__ == "__main__":
    ref = "/root/reference"
    if "--reference" in sys.argv:
        ref = sys.argv[sys.argv.index("--reference") + 1]
    ok = build(ref, force="--force" in sys.argv)
    print("oracle/_ref:", "ok" if ok else "missing")
