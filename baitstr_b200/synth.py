"""ctypes front end of libbaitstr_synth.so (deterministic synthetic FASTQ)."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None

# seed base of SURVEY.md 8(d): 0xBA175712 + config number
SEED_BASE = 0xBA175712


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(os.path.join(HERE, "libbaitstr_synth.so"))
        _lib.synth_record_bytes.argtypes = [C.c_int, C.c_int]
        _lib.synth_record_bytes.restype = C.c_uint64
        _lib.synth_fastq.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_uint64,
                                     C.c_uint64, C.c_int, C.c_uint64]
        _lib.synth_fastq.restype = C.c_uint64
    return _lib


def record_bytes(cfg, L):
    return int(lib().synth_record_bytes(cfg, L))


def generate_into(arr, cfg, L, first, count, mate=1, seed=None):
    """Fill the uint8 numpy array `arr` with reads [first, first+count); returns bytes written."""
    seed = SEED_BASE + cfg if seed is None else seed
    n = lib().synth_fastq(arr.ctypes.data, arr.nbytes, cfg, L, first, count, mate, seed)
    if n == 0 and count:
        raise ValueError("buffer too small for %d reads" % count)
    return int(n)


def generate(cfg, L, first, count, mate=1, seed=None):
    arr = np.empty(record_bytes(cfg, L) * count, dtype=np.uint8)
    n = generate_into(arr, cfg, L, first, count, mate, seed)
    return arr[:n]
