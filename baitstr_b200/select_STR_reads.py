"""Drop-in for the reference's ``select_STR_reads`` extension module.

Same module-level names, argument order and return shapes as
/root/reference/src/select_STR_reads.pyx: ``select_str_reads`` (:339-373, the entry point the
driver script calls), ``MatchPatternsAndReportBest`` (:106), ``ProcessSequence`` (:294),
``ReverseComplement`` (:187), ``PrintSequence`` (:171) and the record type ``fastqsequence``
(src/fastq.pyx:15-19).  The matching happens in libbaitstr_select.so on the GPU; there is no CPU
path here.  ``num_processes`` and ``chunksize`` are accepted and never change the output (they
only size host staging).  ``check_strs`` (``-c FILE``) prints the reference's stderr report
(check_strs.py), ``debug_flag`` (``-d``) the per-read blocks of ProcessSequence (debug_report.py,
enumerated on the GPU by bstr_debug_reads).
"""
import re
import sys

from . import capi
from . import check_strs as check_strs_mod
from . import debug_report

_PATTERN = re.compile(r"^\(\?=\(\(\?=\(\.\{(\d+)\}\)\)\\2\{(\d+),\}\)\)$")      # pyx:350
_COMPLEMENT = bytes.maketrans(b"atcgnATCGN", b"tagcnTAGCN")                      # pyx:188


class fastqsequence(object):
    """src/fastq.pyx:15-19: name (header without '@'), seq, qual."""

    def __init__(self, name, seq, qual):
        self.name, self.seq, self.qual = name, seq, qual


def select_str_reads(filenames, num_minimum_copies, flanking_distance, illumina_quals,
                     check_strs, only_k_mers, remove_homopolymers, debug_flag, altalgo,
                     num_processes, chunksize):
    sys.stdout.flush()
    to_check = check_strs_mod.read_str_list(check_strs)                  # pyx:354, before any read
    ctx = capi.Context(n=num_minimum_copies, f=flanking_distance, ks=only_k_mers,
                       rmhm=remove_homopolymers, altalgo=altalgo, illumina=illumina_quals,
                       host_threads=num_processes)
    try:
        if not check_strs and not debug_flag:
            ctx.run_files(list(filenames), out_fd=1, err_fd=2)
        else:
            _run_files_reported(ctx, list(filenames), to_check if check_strs else None, debug_flag)
    finally:
        ctx.close()
    if check_strs:
        sys.stderr.flush()
        check_strs_mod.print_performance_log(to_check)                   # pyx:372-373
        sys.stderr.flush()


REPORT_BATCH_READS = 20000            # records per GPU batch of a -c / -d run: bounds the memory of the reports


def _run_files_reported(ctx, filenames, to_check, debug_flag, batch_reads=None):
    """``-c`` and/or ``-d``, streamed: the file is read `batch_reads` records at a time; every
    batch goes through the GPU pipeline (bstr_scan_batch), its annotated FASTQ is written to
    stdout unchanged, then its stderr reports follow in the reference's order with one worker:
    per read its -d block (debug_report.py, enumerated on the GPU by bstr_debug_reads) and, if it
    was selected, its -c "No match" line (check_strs.py, from the batch's own output); after the
    last batch "Done processing" (pyx:370).  Memory is bounded by the batch, whatever the file size."""
    import gzip
    import os
    from itertools import islice
    batch_reads = batch_reads or int(os.environ.get("BSTR_REPORT_BATCH_READS", "0")) or REPORT_BATCH_READS
    for fn in filenames:
        opener = gzip.open if fn[-3:] == ".gz" else open                          # fastq.pyx:50-51
        with opener(fn, "rb") as fh:
            ahead = fh.readline()
            rec_base = 0
            while ahead:
                lines = [ahead] + list(islice(fh, 4 * batch_reads - 1))
                ahead = fh.readline()
                raw = b"".join(lines)
                try:
                    out, info = ctx.scan_batch(raw, final=not ahead)
                except capi.BstrError as e:
                    if e.bad_record >= 0:
                        raise capi.BstrError(e.code, "%s: record %d: %s" % (fn, rec_base + e.bad_record, ctx._lib.bstr_last_error(ctx._ctx).decode("latin-1", "replace")), rec_base + e.bad_record)
                    raise
                os.write(1, out)
                check_lines = iter(check_strs_mod.annotated_fastq_lines(out, to_check)) if to_check is not None else None
                if debug_flag:
                    # a selected read's "No match" line follows its own block (both come out of the
                    # same ProcessSequence call, pyx:296-317); names and bases as fastq.pyx:78,83
                    nrec = int(info["n_records"])
                    names = [l[1:-1].decode("latin-1") for l in lines[0:4 * nrec:4]]
                    seqs = [l[:-1].upper() for l in lines[1:4 * nrec:4]]
                    for name, (dinfo, fwd, rc) in zip(names, ctx.debug_reads(seqs)):
                        sys.stderr.write(debug_report.render_read(name, dinfo, fwd, rc))
                        if check_lines is not None and not (dinfo[1] or dinfo[5]):
                            sys.stderr.write(next(check_lines, ""))
                if check_lines is not None:
                    for line in check_lines:
                        sys.stderr.write(line)
                sys.stderr.flush()
                rec_base += int(info["n_records"])
        sys.stderr.write("Done processing %s\n" % fn)                           # pyx:370
        sys.stderr.flush()


def _periods_and_copies(patterns):
    """The reference passes the regex strings built at pyx:350; plain periods are accepted too
    (then the minimum copy number is the reference's default, 2)."""
    ks, n = [], None
    for p in patterns:
        if isinstance(p, int):
            ks.append(p)
            continue
        m = _PATTERN.match(p)
        if not m:
            raise ValueError("not a select_STR_reads pattern: %r" % (p,))
        ks.append(int(m.group(1)))
        if n is not None and n != int(m.group(2)):
            raise ValueError("patterns disagree on the minimum copy number")
        n = int(m.group(2))
    return ks, (2 if n is None else n)


_contexts = {}                       # parameter tuple -> capi.Context (per-read hooks are called in loops)


def _context(n, f, ks, rmhm, altalgo):
    key = (n, f, tuple(sorted(ks)), bool(rmhm), bool(altalgo))
    ctx = _contexts.get(key)
    if ctx is None:
        if len(_contexts) >= 8:
            _contexts.pop(next(iter(_contexts))).close()
        ctx = _contexts[key] = capi.Context(n=n, f=f, ks=list(key[2]), rmhm=key[3], altalgo=key[4])
    return ctx


def _as_bytes(string):
    return string.encode("latin-1") if isinstance(string, str) else bytes(string)


def _tuple(seq, start, ck, as_str):
    """(len(STR), copies, STR, motif, start, end): the tuple the reference collects and
    maximises (pyx:93,101-102 / :149,153)."""
    k, c = ck & 7, ck >> 3
    motif, rep = seq[start:start + k], seq[start:start + k * c]
    if as_str:
        motif, rep = motif.decode("latin-1"), rep.decode("latin-1")
    return (k * c, c, rep, motif, start, start + k * c)


def ReverseComplement(seq):
    """pyx:187-189: complement of acgtn/ACGTN, every other byte unchanged, reversed."""
    if isinstance(seq, str):
        return seq.encode("latin-1").translate(_COMPLEMENT)[::-1].decode("latin-1")
    return bytes(seq).translate(_COMPLEMENT)[::-1]


def MatchPatternsAndReportBest(string, patterns, remove_homopolymers, flanking_distance,
                               debug_flag=False, altalgo=True):
    """pyx:106 (and :58 for altalgo): best (len(STR), copies, STR, motif, start, end) of ONE
    strand, or None.  The string is taken as given (the reference upper-cases in the reader,
    src/fastq.pyx:83, not here; its regex is case-insensitive, the tuple keeps the input's
    case)."""
    ks, n = _periods_and_copies(patterns)
    s = _as_bytes(string)
    ctx = _context(n, flanking_distance, ks, remove_homopolymers, altalgo)
    fstart, _, fck, _ = (int(x) for x in ctx.match_reads([s])[0])
    if fck == 0:
        return None
    return _tuple(s, fstart, fck, isinstance(string, str))


def ProcessSequence(patterns, remove_homopolymers, flanking_distance, debug_flag, altalgo,
                    to_check, check_strs, illumina_quals, s):
    """pyx:294-319: None, or (s, match, rc_match) when both strands have a best STR (rc_match in
    the coordinates of the reverse-complemented sequence)."""
    if s is None:
        return None
    ks, n = _periods_and_copies(patterns)
    seq = _as_bytes(s.seq)
    ctx = _context(n, flanking_distance, ks, remove_homopolymers, altalgo)
    fstart, rstart, fck, rck = (int(x) for x in ctx.match_reads([seq])[0])
    if fck == 0 or rck == 0:
        return None
    as_str = isinstance(s.seq, str)
    return (s, _tuple(seq, fstart, fck, as_str), _tuple(ReverseComplement(seq), rstart, rck, as_str))


def PrintSequence(s, match, rc_match, illumina_quals):
    """pyx:171-185 (host-side formatter of one record; bstr_run_files formats on the GPU)."""
    out = sys.stdout
    out.write("@%s\t%s\t%d\t%d\t%d\t%s\t%d\t%d\t%d\n" % (
        s.name, match[3], match[1], match[4], match[5], rc_match[3], rc_match[1], rc_match[4], rc_match[5]))
    out.write("%s\n+\n" % s.seq)
    out.write("%s\n" % ("".join(chr(ord(x) - 64 + 33) for x in s.qual) if illumina_quals else s.qual))
