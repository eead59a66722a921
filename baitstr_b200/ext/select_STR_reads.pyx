# cython: language_level=3
"""select_STR_reads -- extension module with the reference's module name and entry point.

The reference builds ``select_STR_reads.so`` from src/select_STR_reads.pyx (src/setup.py:4-7) and
its driver script does ``from select_STR_reads import *`` (src/select_STR_reads:61).  This module
keeps that name and the 11-argument ``select_str_reads`` (src/select_STR_reads.pyx:339-341); its
body is a call through the C ABI of libbaitstr_select.so (include/baitstr_select.h) -- the B200
implementation of the stage.  There is no CPU path: without a CUDA device bstr_create fails and
the error is raised.

Runs with the diagnostics of the reference (``check_strs`` / ``debug_flag``, stderr only) and the
per-read hooks (``ProcessSequence``, ``MatchPatternsAndReportBest``, ``ReverseComplement``,
``PrintSequence``) come from the Python package next to the library (baitstr_b200).
"""
import os
import sys

from libc.stdint cimport int32_t, uint64_t
from libc.stdlib cimport malloc, free
from libc.string cimport memset

cdef extern from "baitstr_select.h":
    ctypedef struct bstr_params:
        int32_t abi_version
        int32_t num_copies
        int32_t flank
        int32_t kmask
        int32_t remove_homopolymers
        int32_t altalgo
        int32_t illumina_quals
        int32_t n_devices
        int32_t devices[16]
        int32_t host_threads
        int32_t want_results
        int32_t reserved0
        uint64_t batch_bytes
    ctypedef struct bstr_ctx:
        pass
    int BSTR_ABI_VERSION
    bstr_ctx *bstr_create(const bstr_params *p, char *err, size_t errlen)
    int bstr_run_files(bstr_ctx *ctx, const char *const *paths, int npaths, int out_fd, int err_fd) nogil
    const char *bstr_last_error(const bstr_ctx *ctx)
    void bstr_destroy(bstr_ctx *ctx)


class BstrError(RuntimeError):
    """A negative return code of the C ABI (include/baitstr_select.h) with the library's message."""

    def __init__(self, code, message):
        RuntimeError.__init__(self, "libbaitstr_select error %d: %s" % (code, message))
        self.code = code


def _package():
    """The Python package that holds the diagnostics and the per-read hooks: next to this module
    (installed layout: bin/baitstr_b200) or one directory up (source tree)."""
    here = os.path.dirname(os.path.abspath(__file__))
    for base in (here, os.path.dirname(here)):
        if os.path.isdir(os.path.join(base, "baitstr_b200")) and base not in sys.path:
            sys.path.insert(0, base)
    import baitstr_b200.select_STR_reads as impl
    return impl


def select_str_reads(filenames, num_minimum_copies, flanking_distance, illumina_quals,
                     check_strs, only_k_mers, remove_homopolymers, debug_flag, altalgo,
                     num_processes, chunksize):
    """src/select_STR_reads.pyx:339-373: every file in argv order, selected reads as annotated
    FASTQ on stdout (fd 1), "Done processing <file>" on stderr (fd 2).  num_processes sizes the
    host reader/inflate threads (0: all CPUs), chunksize is accepted; neither changes the output."""
    if check_strs or debug_flag:
        return _package().select_str_reads(filenames, num_minimum_copies, flanking_distance, illumina_quals,
                                           check_strs, only_k_mers, remove_homopolymers, debug_flag, altalgo,
                                           num_processes, chunksize)
    cdef bstr_params p
    cdef char err[512]
    memset(&p, 0, sizeof(p))
    memset(err, 0, sizeof(err))
    p.abi_version = BSTR_ABI_VERSION
    p.num_copies = num_minimum_copies
    p.flank = flanking_distance
    for k in only_k_mers:
        if 2 <= int(k) <= 6:
            p.kmask |= 1 << int(k)
    p.remove_homopolymers = 1 if remove_homopolymers else 0
    p.altalgo = 1 if altalgo else 0
    p.illumina_quals = 1 if illumina_quals else 0
    p.host_threads = num_processes if num_processes and num_processes > 0 else 0
    names = [os.fsencode(f) for f in filenames]
    cdef int n = len(names)
    cdef bstr_ctx *ctx = bstr_create(&p, err, sizeof(err))
    if ctx == NULL:
        raise BstrError(-1, (<bytes>err).decode("latin-1", "replace"))
    cdef const char **paths = <const char **>malloc((n + 1) * sizeof(char *))
    cdef int i, rc
    for i in range(n):
        paths[i] = names[i]
    sys.stdout.flush()
    sys.stderr.flush()
    with nogil:
        rc = bstr_run_files(ctx, paths, n, 1, 2)
    message = (<bytes>bstr_last_error(ctx)).decode("latin-1", "replace")
    free(paths)
    bstr_destroy(ctx)
    if rc:
        raise BstrError(rc, message)


def MatchPatternsAndReportBest(*args, **kw):
    return _package().MatchPatternsAndReportBest(*args, **kw)


def ProcessSequence(*args, **kw):
    return _package().ProcessSequence(*args, **kw)


def ReverseComplement(seq):
    return _package().ReverseComplement(seq)


def PrintSequence(s, match, rc_match, illumina_quals):
    return _package().PrintSequence(s, match, rc_match, illumina_quals)
