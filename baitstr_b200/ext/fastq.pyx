# cython: language_level=3
"""fastq -- import-compatible stand-in for the reference's FASTQ reader module.

The reference's select stage reads its input through ``fastq.fastq`` (src/fastq.pyx:48-94); in this
build the records are split, checked and upper-cased on the GPU (csrc/fused.cuh), so nothing on
the select path imports this module.  It exists because the installed layout of the reference
has a ``fastq.so`` (Makefile:6) that other scripts may import: same class names, same record
fields, same acceptance rules (header '@', bases in acgtnACGTN., separator '+', ".gz" by file
name suffix, the last character of every line dropped), written for Python 3 text or bytes.
"""
import gzip

_BASES = frozenset(b"acgtnACGTN.")
_COMPLEMENT = bytes.maketrans(b"atcgnATCGN", b"tagcnTAGCN")


class fastqsequence(object):
    """One record: name (header without '@'), seq (upper case), qual."""

    def __init__(self, name, seq, qual):
        self.name, self.seq, self.qual = name, seq, qual

    def __len__(self):
        if len(self.seq) != len(self.qual):
            raise AssertionError("sequence and quality lengths differ")
        return len(self.seq)

    def __str__(self):
        return "@%s\n%s\n+\n%s" % (self.name, self.seq, self.qual)

    def fq2seqfa(self):
        return ">%s\n%s" % (self.name, self.seq)

    def fq2qfa(self, offset):
        return ">%s\n%s" % (self.name, " ".join(str(ord(q) - offset) for q in self.qual))

    def reverse_complement(self):
        self.seq = self.seq.encode("latin-1").translate(_COMPLEMENT)[::-1].decode("latin-1")
        self.qual = self.qual[::-1]


class fastq(object):
    """Iterator over the records of a FASTQ file; ``records.fastqsequence`` is the current one
    (the reference yields the reader itself, src/fastq.pyx:92-94)."""

    def __init__(self, filename):
        name = filename.decode("latin-1") if isinstance(filename, bytes) else filename
        opener = gzip.open if name.endswith(".gz") else open
        self.file = opener(name, "rb")
        self.fastqsequence = None

    def __iter__(self):
        return self

    def close(self):
        self.file.close()

    @staticmethod
    def checkbases(line):
        body = line[:-1]
        if isinstance(body, str):
            body = body.encode("latin-1")
        return _BASES.issuperset(body)

    def __next__(self):
        header = self.file.readline()
        if not header:
            self.file.close()
            raise StopIteration
        if header[:1] != b"@":
            raise AssertionError("header should begin with a @ in a fastq file")
        bases = self.file.readline()
        if not fastq.checkbases(bases):
            raise AssertionError("read %s should only have ACGTN" % header[1:-1].decode("latin-1"))
        plus = self.file.readline()
        if plus[:1] != b"+":
            raise AssertionError("separator line in fastq file should begin with a +")
        qual = self.file.readline()
        self.fastqsequence = fastqsequence(header[1:-1].decode("latin-1"), bases[:-1].upper().decode("latin-1"),
                                           qual[:-1].decode("latin-1"))
        return self

    next = __next__
