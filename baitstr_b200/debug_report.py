"""``-d/--debug``: the per-read stderr report of the reference's ProcessSequence
(/root/reference/src/select_STR_reads.pyx:294-311) with the messages of its two matchers
(:76-87 default algorithm, :138-140 and :157-163 ``-x``).

The regex hits and their verdicts are enumerated on the GPU (``bstr_debug_reads``,
csrc/matcher.cuh: debug_read); this module only turns the codes into the reference's lines.
stdout is never affected.  With more than one worker process the reference's blocks interleave
nondeterministically; the order here is the input order (= the reference with ``-p 1``).
"""
import gzip

RULE = "-" * 79                                                                   # pyx:297
_LINES = {ord("N"): "N in motif.", ord("5"): "Insufficient 5' flank.", ord("3"): "Insufficient 3' flank."}


def render_read(name, info, fwd_codes, rc_codes):
    """One read's block.  info = 8 integers as written by bstr_debug_reads."""
    out = [RULE, name]
    for strand, codes in ((0, fwd_codes), (1, rc_codes)):
        count, is_none, tail_code, tail_value = (int(x) for x in info[4 * strand:4 * strand + 4])
        out.extend(_LINES[c] for c in bytes(codes[:count]))
        if tail_code:                                                             # pyx:157-163
            out.append("Insufficient %s' flank.: %d" % (chr(tail_code), tail_value))
        if is_none:
            out.append("0 valid matches.")                                        # pyx:303-304, 310-311
            break
    return "\n".join(out) + "\n"


def read_names_and_bases(path):
    """(name, upper-cased bases) of every record of a FASTQ file the GPU pipeline has already
    accepted (src/fastq.pyx:78,83: header without '@', bases upper-cased)."""
    opener = gzip.open if path[-3:] == ".gz" else open                            # fastq.pyx:50-51
    with opener(path, "rb") as fh:
        while True:
            header = fh.readline()
            if not header:
                return
            bases = fh.readline()
            fh.readline()
            fh.readline()
            yield header[1:-1].decode("latin-1"), bases[:-1].upper()


def file_blocks(ctx, path, batch_reads=20000):
    """Yields (block text, read is selected) for every read of `path`, in input order."""
    names, seqs = [], []

    def flush():
        for name, (info, fwd, rc) in zip(names, ctx.debug_reads(seqs)):
            yield render_read(name, info, fwd, rc), not (info[1] or info[5])
        del names[:], seqs[:]

    for name, bases in read_names_and_bases(path):
        names.append(name)
        seqs.append(bases)
        if len(names) >= batch_reads:
            yield from flush()
    if names:
        yield from flush()


def report_file(ctx, path, err, batch_reads=20000):
    """Write the blocks of every read of `path` to the text stream `err`."""
    for block, _selected in file_blocks(ctx, path, batch_reads):
        err.write(block)
