#!/usr/bin/env python3
"""Run the UNMODIFIED reference select stage (built by oracle/build_ref.py) under Python 3.

TEST INFRASTRUCTURE / CPU BASELINE ONLY.  Nothing in the product path imports this.

The reference driver script (/root/reference/src/select_STR_reads:69-140) is Python 2
and cannot run here; this runner applies the four runtime shims of SURVEY.md 8(c) and
repeats the driver's option handling with the same getopt spec (select_STR_reads:74),
then calls the reference's own select_str_reads(...) (select_STR_reads.pyx:339).

  shim 1  itertools.izip_longest / string.maketrans      (select_STR_reads.pyx:3-4)
  shim 2  fastq.fastq.__next__ = fastq.fastq.next         (fastq.pyx:72, Py2 iterator)
  shim 3  module-global `fastq` replaced by a line-for-line Py3 restatement of
          fastq.pyx:48-94 that opens latin-1 text (8-bit clean) and honours ".gz"
          (under c_string_type=str the original compares str to bytes at fastq.pyx:50
          and silently takes the plain-file branch)
  shim 4  this driver itself
"""
import gzip
import itertools
import os
import string
import sys
from getopt import getopt, GetoptError

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")


def load_reference():
    """Import the reference extension modules with the shims applied."""
    if not hasattr(itertools, "izip_longest"):
        itertools.izip_longest = itertools.zip_longest
    if not hasattr(string, "maketrans"):
        string.maketrans = str.maketrans
    if REF not in sys.path:
        sys.path.insert(0, REF)
    import fastq as ref_fastq                 # noqa: E402  (oracle/_ref/fastq.*.so)
    import select_STR_reads as ref_select     # noqa: E402
    ref_fastq.fastq.__next__ = ref_fastq.fastq.next
    fastqsequence = ref_fastq.fastqsequence

    class PyFastq:
        """fastq.pyx:48-94 restated for Python 3 text I/O (same asserts, same slicing)."""

        def __init__(self, filename):
            if filename[-3:] == ".gz":
                self.file = gzip.open(filename, "rt", encoding="latin-1", newline="\n")
            else:
                self.file = open(filename, "r", encoding="latin-1", newline="\n")
            self.fastqsequence = None

        def __iter__(self):
            return self

        @staticmethod
        def checkbases(line):
            for base in line[:-1]:
                if base not in ["a", "c", "g", "t", "n", "A", "C", "G", "T", "N", "."]:
                    return False
            return True

        def close(self):
            self.file.close()

        def __next__(self):
            line = self.file.readline()
            if not line:
                self.file.close()
                raise StopIteration
            assert line[0] == "@", "header should begin with a @ in a fastq file"
            name = line[1:-1]
            line = self.file.readline()
            assert PyFastq.checkbases(line) is True, "read %s should only have ACGTN" % name
            seq = line[:-1].upper()
            line = self.file.readline()
            assert line[0] == "+", "separator line in fastq file should begin with a +"
            line = self.file.readline()
            qual = line[:-1]
            self.fastqsequence = fastqsequence(name, seq, qual)
            return self

        next = __next__

    ref_select.fastq = PyFastq
    return ref_select


USAGE = "usage: ref_select.py [options] reads1.fq reads2.fq ...  (options as select_STR_reads)"


def main(argv):
    ref_select = load_reference()
    with open(os.path.join(REF, "VERSION")) as fh:
        version = fh.read().strip()
    try:
        opts, args = getopt(argv, "hdn:f:ic:v23456rxp:s:",
                            ["help", "debug", "numcopies=", "flanks=", "illumina", "check=",
                             "version", "2mer", "3mer", "4mer", "5mer", "6mer", "removehm",
                             "altalgo", "numprocesses=", "chunksize="])
    except GetoptError as err:
        print(str(err))
        print(USAGE, file=sys.stderr)
        return 2
    n, f, illumina, check, print_version = 2, 31, False, None, False
    ks, rmhm, debug, altalgo, nproc, chunk = [], False, False, True, 2, 100000
    for o, a in opts:
        if o in ("-h", "--help"):
            print(USAGE, file=sys.stderr)
            return 0
        elif o in ("-d", "--debug"):
            debug = True
        elif o in ("-n", "--numcopies"):
            n = int(a)
        elif o in ("-f", "--flanks"):
            f = int(a)
        elif o in ("-i", "--illumina"):
            illumina = True
        elif o in ("-c", "--check"):
            check = a
        elif o in ("-v", "--version"):
            print_version = True
        elif o in ("-2", "-3", "-4", "-5", "-6"):
            ks.append(int(o[1]))
        elif o in ("--2mer", "--3mer", "--4mer", "--5mer", "--6mer"):
            ks.append(int(o[2]))
        elif o in ("-r", "--removehm"):
            rmhm = True
        elif o in ("-x", "--altalgo"):
            altalgo = False
        elif o in ("-p", "--numprocesses"):
            nproc = int(a)
        elif o in ("-s", "--chunksize"):
            chunk = int(a)
    if print_version:
        print("Program: select_STR_reads", file=sys.stderr)
        print("Version: %s" % version, file=sys.stderr)
        print("Contact: Aakrosh Ratan (ratan@bx.psu.edu)", file=sys.stderr)
        return 0
    if len(args) < 1:
        print(USAGE, file=sys.stderr)
        return 3
    ks = list(set(ks))
    if len(ks) == 0:
        ks = [2, 3, 4, 5, 6]
    ref_select.select_str_reads(args, n, f, illumina, check, ks, rmhm, debug, altalgo,
                                nproc, chunk)
    sys.stdout.flush()
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
