"""Drop-in for the reference's ``select_STR_reads`` extension module.

Same entry point, same 11 positional arguments, same stdout/stderr behaviour as
/root/reference/src/select_STR_reads.pyx:339-373; the work happens in
libbaitstr_select.so on the GPU (bstr_run_files).  ``num_processes`` and ``chunksize`` are
accepted and never change the output (they only size host staging).
"""
import sys

from . import capi


def select_str_reads(filenames, num_minimum_copies, flanking_distance, illumina_quals,
                     check_strs, only_k_mers, remove_homopolymers, debug_flag, altalgo,
                     num_processes, chunksize):
    sys.stdout.flush()
    ctx = capi.Context(n=num_minimum_copies, f=flanking_distance, ks=only_k_mers,
                       rmhm=remove_homopolymers, altalgo=altalgo, illumina=illumina_quals,
                       host_threads=num_processes)
    try:
        ctx.run_files(list(filenames), out_fd=1, err_fd=2)
    finally:
        ctx.close()


def MatchPatternsAndReportBest(string, only_k_mers, remove_homopolymers, flanking_distance,
                               num_minimum_copies=2, altalgo=True):
    """Per-strand test hook, the GPU counterpart of select_STR_reads.pyx:106 (batch of one).
    Returns (motif, copies, start, end) or None for the string as given (one strand)."""
    with capi.Context(n=num_minimum_copies, f=flanking_distance, ks=only_k_mers,
                      rmhm=remove_homopolymers, altalgo=altalgo) as ctx:
        s = string.encode("latin-1") if isinstance(string, str) else bytes(string)
        fstart, _, fck, _ = (int(x) for x in ctx.match_reads([s])[0])
    if fck == 0:
        return None
    k, c = fck & 7, fck >> 3
    return (s[fstart:fstart + k].decode("latin-1").upper(), c, fstart, fstart + k * c)
