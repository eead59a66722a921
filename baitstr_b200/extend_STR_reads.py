"""Host side of the extend stage: the command line of the reference's extend_STR_reads
(/root/reference/src/extend_STR_reads.c:1008-1119 with the option parser of clparsing.c) over
bstr_kmers_* and bstr_kmers_extend (include/baitstr_select.h), which count the k-mers and walk the
blocks on the GPU.  Same options, same usage text, same stdout; stderr carries the reference's main
messages.  There is no CPU implementation: without libbaitstr_select.so and a CUDA device it fails."""
import ctypes as C
import gzip
import sys
import time

from . import capi
from .merge_STR_reads import Die, as_bool, as_uint, parse, usage

PROGRAM = ("extend_STR_reads", "Extend fastq reads based on the kmer structure",
           "extend_STR_reads [options] gs cov klen str.reads.fq reads1.fq reads2.fq ...")
OPTIONS = [
    ("help", "FALSE", False, True, "Print this string and quit."),
    ("debug", "FALSE", False, True, "Print extra debug information in this run."),
    ("memDebug", "FALSE", False, False, "Print memory related debug information in this run."),
    ("min_threshold", "2", True, True, "Discard kmers that are observed < min_threshold"),
    ("max_threshold", "255", True, True, "Discard kmers that are observed > max_threshold"),
    ("progress", "1000000", True, True, "print progress every so many sequences"),
    ("flanks", "1024", True, True, "the maximum size of flanks extension"),
    ("ploidy", "2", True, True, "the ploidy of the genome"),
    ("heterozygosity", "0.001", True, True, "fraction of nucleotides that differ between inherited chromosomes"),
    ("errorrate", "0.01", True, True, "expected error rate in sequencing"),
]
CHUNK = 1 << 30         # bytes of reads per bstr_kmers_add_fastq call (whole records)


def as_double(values, name):
    try:
        return float(values[name])
    except ValueError:
        raise Die("%s: Expected an integer, but did not find it\n" % name)      # the reference's wording (clparsing.c:153-156)


def read_file(path):
    with open(path, "rb") as f:
        raw = f.read()
    if raw[:2] == b"\x1f\x8b":          # gzopen reads either (fastq_seq.c:9)
        raw = gzip.decompress(raw)
    return raw          # an unterminated last line is an error below, as in the reference's reader (utilities.c:205)


def record_chunks(raw, limit=CHUNK):
    """pieces of at most `limit` bytes that end on a record boundary (every 4th newline)"""
    pos = 0
    while pos < len(raw):
        if len(raw) - pos <= limit:
            yield raw[pos:]
            return
        lines = raw.count(b"\n", pos, pos + limit)
        cut = raw.rfind(b"\n", pos, pos + limit) + 1           # behind the last newline of the window ...
        for _ in range(lines % 4):                               # ... and back over the lines of the record it cuts
            cut = raw.rfind(b"\n", pos, cut - 1) + 1
        if lines < 4:
            cut = pos
        if cut == pos:
            raise Die("a record longer than %d bytes" % limit)
        yield raw[pos:cut]
        pos = cut


def main(argv, version, stdout=None, stderr=None):
    stdout = stdout or sys.stdout
    stderr = stderr or sys.stderr
    t0 = time.time()

    def note(text):                     # PrintDebugMessage: "[%d s]\t..." on stderr
        stderr.write("[%d s]\t%s\n" % (int(time.time() - t0), text))

    try:
        values, args = parse(argv, OPTIONS)
        if as_bool(values, "help"):
            stdout.write(usage(version, PROGRAM, OPTIONS))
            return 0
        if len(args) < 6:               # argc < 7 (:1046-1049): two read files at least
            stdout.write(usage(version, PROGRAM, OPTIONS))
            return 1
        for what, a in (("Expected genome size should be an integer", args[0]), ("Expected coverage should be an integer", args[1]),
                        ("Kmer length should be an odd integer < 32", args[2])):
            if not a.lstrip()[:1].isdigit():
                raise Die("%s: %s" % (what, a))
        gs, cov, klen = as_uint({"v": args[0]}, "v"), as_uint({"v": args[1]}, "v"), as_uint({"v": args[2]}, "v")
        if klen % 2 == 0:
            klen -= 1
            stderr.write("Kmer length should be an odd integer, using %d\n" % klen)
        if not 2 <= klen <= 31:
            raise Die("Kmer length should be an odd integer < 32: %s" % args[2])
        lo, hi, flanks = as_uint(values, "min_threshold"), as_uint(values, "max_threshold"), as_uint(values, "flanks")
        ploidy, het, err = as_uint(values, "ploidy"), as_double(values, "heterozygosity"), as_double(values, "errorrate")
        as_uint(values, "progress")
        if as_bool(values, "debug"):
            raise Die("--debug (the reference's per-k-mer trace) is not available in the GPU build")
        genome = int(gs * (1 + het * (ploidy - 1) * klen))                                  # :868-869
        expected = int(genome * (1 + (cov * (1 - pow(1 - err, klen)))))
        note("Expecting %d kmers in this dataset with haploid genome size %d bps.\n" % (expected, gs))
        blocks = read_file(args[3])
        files = [blocks] + [read_file(p) for p in args[4:]]
        # the table must hold every distinct k-mer, singletons included: no more than the k-mers there are
        upper = sum(max(0, len(f) // 2 - f.count(b"\n") // 4 * (klen - 1)) for f in files)
        lib = capi.load_library()
        with capi.Context() as ctx:
            h = lib.bstr_kmers_create(ctx._ctx, klen, max(1024, min(upper, 1 << 31)))
            if not h:
                ctx._raise(-3)
            try:
                for path, raw in zip(args[3:], files):          # the reference counts argv[4 ..]: the blocks' file first (:890-897)
                    for piece in record_chunks(raw):
                        rc = lib.bstr_kmers_add_fastq(h, piece, len(piece))
                        if rc:
                            ctx._raise(rc)
                    note("1. Done with all the sequences in %s" % path)
                st = (C.c_uint64 * 4)()
                stdout.flush()
                fd = getattr(stdout, "fileno", lambda: -1)()
                import tempfile
                with tempfile.TemporaryFile() as f:
                    rc = lib.bstr_kmers_extend(h, min(lo, 0x7FFFFFFF), min(hi, 0x7FFFFFFF), min(flanks, 0x7FFFFFFF), blocks, len(blocks), f.fileno(), st)
                    if rc:
                        ctx._raise(rc)
                    f.seek(0)
                    out = f.read()
            finally:
                lib.bstr_kmers_destroy(h)
        note("Read %d kmers that are observed at least 2 times." % st[3])
        note("Done with all extensions in %s" % args[3])
        if st[2]:                       # IssueWarningAboutKmerLength + the closing count (:853-861, :1000-1006)
            stderr.write("\n==========================================================================\n"
                         "%d ends of merged reads were not extended as the kmer length for extension \n"
                         "is greater than the flank requirement used for STR discovery in select_STR_reads\n"
                         "\n==========================================================================\n" % st[2])
        getattr(stdout, "buffer", stdout).write(out)
        stdout.flush()
        return 0
    except Die as e:
        stderr.write("%s\n" % e)
        return 1
    except capi.BstrError as e:
        stderr.write("extend_STR_reads: %s\n" % e)
        return 1
    except OSError as e:
        stderr.write("extend_STR_reads: %s\n" % e)
        return 1
