#!/usr/bin/env python3
"""Build the two extension modules of the installed layout -- select_STR_reads.so and fastq.so,
the names the reference installs (/root/reference/Makefile:6-7, src/setup.py:4-7) -- into DEST:
cython -> gcc, select_STR_reads linked against libbaitstr_select.so found next to it at run time
(rpath $ORIGIN).  Usage: python build_ext.py DEST_DIR LIB_DIR"""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))


def build(dest, libdir):
    os.makedirs(dest, exist_ok=True)
    tmp = os.path.join(dest, ".ext_build")
    os.makedirs(tmp, exist_ok=True)
    inc = sysconfig.get_paths()["include"]
    for mod, extra in (("fastq", []),
                       ("select_STR_reads", ["-I", os.path.join(ROOT, "include"), "-L", libdir, "-lbaitstr_select",
                                             "-Wl,-rpath,$ORIGIN", "-Wl,-rpath," + os.path.abspath(libdir)])):
        c = os.path.join(tmp, mod + ".c")
        subprocess.check_call([sys.executable, "-m", "cython", "-3", "-o", c, os.path.join(HERE, mod + ".pyx")])
        subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-w", "-fwrapv", "-I", inc, c,
                               "-o", os.path.join(dest, mod + ".so")] + extra)
    for f in os.listdir(tmp):
        os.unlink(os.path.join(tmp, f))
    os.rmdir(tmp)


if __name__ == "__main__":
    build(sys.argv[1], sys.argv[2])
