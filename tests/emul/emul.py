"""ctypes loader for tests/emul/emul_matcher.cpp (host build of the DEVICE matcher source).

TEST INFRASTRUCTURE: lets the GPU-less container fuzz baitstr_b200/csrc/matcher.cuh against
the oracle.  Nothing in the product imports this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
SRC = os.path.join(HERE, "emul_matcher.cpp")
HDRS = [os.path.join(ROOT, "baitstr_b200", "csrc", h) for h in ("matcher.cuh",)]
LIB = os.path.join(HERE, "_build", "libemul_matcher.so")


class Params(C.Structure):
    _fields_ = [("n", C.c_int), ("f", C.c_int), ("kmask", C.c_int), ("rmhm", C.c_int),
                ("altalgo", C.c_int)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        newest = max(os.path.getmtime(p) for p in [SRC] + HDRS)
        if not os.path.exists(LIB) or os.path.getmtime(LIB) < newest:
            os.makedirs(os.path.dirname(LIB), exist_ok=True)
            subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-w", "-o", LIB, SRC])
        _lib = C.CDLL(LIB)
        _lib.emul_match_reads.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(Params),
                                          C.c_int, C.c_void_p]
        _lib.emul_match_reads.restype = None
        _lib.emul_bases_invalid.argtypes = [C.c_char_p, C.c_int]
        _lib.emul_bases_invalid.restype = C.c_int
    return _lib


def match_reads(seqs, n=2, f=31, ks=None, rmhm=False, altalgo=True, force_w=0):
    """seqs: list of bytes -> uint32 array [nreads, 4] = (fstart, rstart, fck, rck)."""
    km = 0
    for k in (ks or [2, 3, 4, 5, 6]):
        km |= 1 << k
    off = np.zeros(len(seqs) + 1, dtype=np.uint64)
    off[1:] = np.cumsum([len(s) for s in seqs])
    blob = np.frombuffer(b"".join(seqs) + b"\0" * 8, dtype=np.uint8).copy()
    res = np.zeros((len(seqs), 4), dtype=np.uint32)
    p = Params(n, f, km, int(rmhm), int(altalgo))
    lib().emul_match_reads(blob.ctypes.data, off.ctypes.data, len(seqs), C.byref(p), force_w,
                           res.ctypes.data)
    return res


def debug_read(seq, n=2, f=31, ks=None, rmhm=False, altalgo=True):
    """bytes -> (info[8], forward codes, reverse-strand codes) through matcher.cuh: debug_read."""
    km = 0
    for k in (ks or [2, 3, 4, 5, 6]):
        km |= 1 << k
    L = lib()
    L.emul_debug_codes_cap.restype = C.c_uint32
    cap = L.emul_debug_codes_cap(len(seq), n, km)
    fwd = np.zeros(cap + 1, dtype=np.uint8)
    rc = np.zeros(cap + 1, dtype=np.uint8)
    info = np.zeros(8, dtype=np.uint32)
    p = Params(n, f, km, int(rmhm), int(altalgo))
    L.emul_debug_read(bytes(seq), len(seq), C.byref(p), fwd.ctypes.data_as(C.c_void_p),
                      rc.ctypes.data_as(C.c_void_p), info.ctypes.data_as(C.c_void_p))
    return info, fwd[:int(info[0])], rc[:int(info[4])]


def bases_invalid(data, misalign=0):
    buf = C.create_string_buffer(b"\0" * misalign + bytes(data))
    return bool(lib().emul_bases_invalid(C.cast(C.addressof(buf) + misalign, C.c_char_p), len(data)))
