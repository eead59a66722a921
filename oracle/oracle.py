"""ctypes front end of oracle/str_oracle.c -- TEST INFRASTRUCTURE, never a product path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg import this.
"""
import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "libstr_oracle.so")


def build(force=False):
    src = os.path.join(HERE, "str_oracle.c")
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(LIB), exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-Wall", "-o", LIB, src])
    return LIB


class Params(C.Structure):
    _fields_ = [("n", C.c_int), ("f", C.c_int), ("kmask", C.c_int), ("rmhm", C.c_int),
                ("altalgo", C.c_int), ("illumina", C.c_int)]


class Match(C.Structure):
    _fields_ = [("found", C.c_int), ("k", C.c_int), ("copies", C.c_int),
                ("start", C.c_int), ("end", C.c_int)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.oracle_match_strand.argtypes = [C.c_char_p, C.c_int, C.POINTER(Params), C.POINTER(Match)]
        _lib.oracle_match_strand.restype = None
        _lib.oracle_revcomp.argtypes = [C.c_char_p, C.c_int, C.c_char_p]
        _lib.oracle_select_buffer.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(Params),
                                              C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t,
                                              C.POINTER(C.c_long), C.POINTER(C.c_long),
                                              C.POINTER(C.c_long)]
        _lib.oracle_select_buffer.restype = C.c_long
    return _lib


def kmask_of(ks):
    m = 0
    for k in (ks or [2, 3, 4, 5, 6]):
        m |= 1 << int(k)
    return m


def make_params(n=2, f=31, ks=None, rmhm=False, altalgo=True, illumina=False):
    return Params(n, f, kmask_of(ks), int(rmhm), int(altalgo), int(illumina))


def match_strand(seq, n=2, f=0, ks=None, rmhm=False, altalgo=True):
    """Best match of one (upper-cased) strand string -> (motif, copies, start, end) or None."""
    b = seq.encode("latin-1") if isinstance(seq, str) else bytes(seq)
    p, m = make_params(n, f, ks, rmhm, altalgo), Match()
    lib().oracle_match_strand(b, len(b), C.byref(p), C.byref(m))
    if not m.found:
        return None
    return (b[m.start:m.start + m.k].decode("latin-1"), m.copies, m.start, m.end)


def revcomp(seq):
    b = seq.encode("latin-1") if isinstance(seq, str) else bytes(seq)
    out = C.create_string_buffer(len(b) + 1)
    lib().oracle_revcomp(b, len(b), out)
    return out.raw[:len(b)].decode("latin-1")


class OracleError(Exception):
    def __init__(self, code, record):
        super().__init__("oracle error %d at record %d" % (code, record))
        self.code, self.record = code, record


def select_buffer(raw, n=2, f=31, ks=None, rmhm=False, altalgo=True, illumina=False,
                  want_results=False):
    """Whole select stage on a FASTQ file image -> output bytes (and per-record results)."""
    import numpy as np
    raw = bytes(raw)
    p = make_params(n, f, ks, rmhm, altalgo, illumina)
    nlines = raw.count(b"\n") + 1
    nrec_cap = nlines // 4 + 2
    outcap = len(raw) + 160 * nrec_cap + 1024
    out = np.empty(outcap, dtype=np.uint8)
    res = np.zeros((nrec_cap, 8), dtype=np.int32)
    nrec, nsel, bad = C.c_long(), C.c_long(), C.c_long()
    rv = lib().oracle_select_buffer(raw, len(raw), C.byref(p), out.ctypes.data, outcap,
                                    res.ctypes.data, nrec_cap, C.byref(nrec), C.byref(nsel),
                                    C.byref(bad))
    if rv < 0:
        raise OracleError(rv, bad.value)
    data = out[:rv].tobytes()
    if want_results:
        return data, res[:nrec.value].copy()
    return data
