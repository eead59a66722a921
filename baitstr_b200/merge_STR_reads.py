"""Host side of the merge stage: the command line of the reference's merge_STR_reads
(/root/reference/src/merge_STR_reads.c:512-588 with the option parser of clparsing.c:205-316) over
bstr_merge_reads (include/baitstr_select.h), which does the work on the GPU.  Same options, same
usage text, same stdout; stderr carries the reference's progress lines (its elapsed-seconds prefix
included).  There is no CPU implementation: without libbaitstr_select.so and a CUDA device it fails."""
import gzip
import sys
import time

from . import capi

PROGRAM = ("merge_STR_reads", "Merge reads that support the same STR", "merge_STR_reads [options] klength reads.str.fq")
# clparsing.c:8-44 (help, debug; memDebug undocumented) + merge_STR_reads.c:524-531, in the order the
# reference lists them
OPTIONS = [
    # name, default, takes a value, documented, help
    ("help", "FALSE", False, True, "Print this string and quit."),
    ("debug", "FALSE", False, True, "Print extra debug information in this run."),
    ("memDebug", "FALSE", False, False, "Print memory related debug information in this run."),
    ("min_threshold", "4", True, True, "Discard blocks that include < min_threshold reads"),
    ("max_threshold", "10000", True, True, "Discard blocks that include > max_threshold reads"),
    ("progress", "1000000", True, True, "print progress every so many sequences"),
    ("all", "FALSE", False, True, "include non-polymorphic blocks"),
]


def usage(version, program=None, options=None):
    """PrintSimpleUsageString, clparsing.c:268-308"""
    program, options = program or PROGRAM, options or OPTIONS
    out = ["Program : %s (%s)\n" % program[:2], "Version : %s\n\n" % version, "usage:\n", "\t%s\n\n" % program[2], "options:\n"]
    width = max(len(o[0]) for o in options if o[3])
    for name, default, valued, documented, text in options:
        if not documented:
            continue
        tail = "[--%s=%s]" % (name, default) if valued else ("[--%s]" % name if default == "TRUE" else "[--no%s]" % name)
        out.append("\t%s%s  %s%s\n" % (name, " " * (width - len(name)), text, tail))
    return "".join(out)


class Die(Exception):
    pass


def parse(argv, options=None):
    """ParseOptions, clparsing.c:205-263: --name, --noname, --name=value anywhere; the rest are arguments"""
    values = dict((o[0], o[1]) for o in (options or OPTIONS))
    rest = []
    for a in argv:
        if a.startswith("--"):
            if "=" in a:
                name, value = a[2:].split("=", 1)
            elif a.startswith("--no"):
                name, value = a[4:], "FALSE"
            else:
                name, value = a[2:], "TRUE"
            if name not in values:
                raise Die("unknown option %s" % name)
            values[name] = value
        else:
            rest.append(a)
    return values, rest


def as_bool(values, name):
    if values[name] not in ("TRUE", "FALSE"):
        raise Die("%s: expected TRUE or FALSE" % name)
    return values[name] == "TRUE"


def as_uint(values, name):
    """sscanf("%u"): leading digits"""
    s = values[name].lstrip()
    digits = ""
    for c in s[1:] if s[:1] == "+" else s:
        if not c.isdigit():
            break
        digits += c
    if not digits:
        raise Die("%s: Expected unsigned int, but did not find it\n" % name)
    return int(digits) & 0xFFFFFFFF


def main(argv, version, stdout=None, stderr=None):
    stdout = stdout or sys.stdout
    stderr = stderr or sys.stderr
    t0 = time.time()

    def note(text):                 # PrintDebugMessage: "[%d s]\t..." on stderr
        stderr.write("[%d s]\t%s\n" % (int(time.time() - t0), text))

    try:
        values, args = parse(argv)
        if as_bool(values, "help"):
            stdout.write(usage(version))
            return 0
        if len(args) < 2:           # the reference tests argc < 2 and then reads argv[2]; here both arguments are required
            stdout.write(usage(version))
            return 1
        klength = as_uint({"k": args[0]}, "k") if args[0].lstrip()[:1].isdigit() else None
        if klength is None:
            raise Die("Kmer length should be an odd integer < 32: %s" % args[0])
        if klength % 2 == 0:
            klength -= 1
            stderr.write("Kmer length should be an odd integer, using %d\n" % klength)
        lo, hi, include_all = as_uint(values, "min_threshold"), as_uint(values, "max_threshold"), as_bool(values, "all")
        progress = as_uint(values, "progress")
        if as_bool(values, "debug"):
            raise Die("--debug (the reference's per-read trace) is not available in the GPU build")
        with open(args[1], "rb") as f:
            raw = f.read()
        if raw[:2] == b"\x1f\x8b":  # gzopen reads either (fastq_seq.c:9)
            raw = gzip.decompress(raw)
        nreads = raw.count(b"\n") // 4
        if progress and nreads:     # merge_STR_reads.c:341-346: reads 1, 1 + progress, ...
            pos = 0
            for i in range(0, nreads, progress):
                if i:
                    for _ in range(4 * progress):
                        pos = raw.index(b"\n", pos) + 1
                note("Processing read number %d: %s" % (i + 1, raw[pos + 1:raw.index(b"\n", pos)].decode("latin-1")))
        with capi.Context() as ctx:
            out, _ = ctx.merge_reads(raw, klength, min(lo, 0x7FFFFFFF), min(hi, 0x7FFFFFFF), include_all)
        note("Processed %d reads." % nreads)
        stdout.flush()
        getattr(stdout, "buffer", stdout).write(out)
        stdout.flush()
        return 0
    except Die as e:
        stderr.write("%s\n" % e)
        return 1
    except capi.BstrError as e:
        stderr.write("merge_STR_reads: %s\n" % e)
        return 1
    except OSError as e:
        stderr.write("merge_STR_reads: %s\n" % e)
        return 1
