"""ctypes binding of include/baitstr_select.h (the drop-in C ABI).

Every call goes to libbaitstr_select.so; if the library is missing or there is no CUDA
device the constructors raise -- there is no fallback path.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ABI_VERSION = 2
NOALLOC_MAX_BYTES = 4 << 20           # BSTR_NOALLOC_MAX_BYTES


def library_path():
    # BSTR_LIB: another build of the same library (kernel A/B measurements only)
    if os.environ.get("BSTR_LIB"):
        return os.environ["BSTR_LIB"]
    here = os.path.join(HERE, "libbaitstr_select.so")
    up = os.path.join(os.path.dirname(HERE), "libbaitstr_select.so")     # installed layout: bin/libbaitstr_select.so
    return here if os.path.exists(here) or not os.path.exists(up) else up


class Params(C.Structure):
    _fields_ = [("abi_version", C.c_int32), ("num_copies", C.c_int32), ("flank", C.c_int32),
                ("kmask", C.c_int32), ("remove_homopolymers", C.c_int32), ("altalgo", C.c_int32),
                ("illumina_quals", C.c_int32), ("n_devices", C.c_int32),
                ("devices", C.c_int32 * 16), ("host_threads", C.c_int32),
                ("want_results", C.c_int32), ("reserved0", C.c_int32),
                ("batch_bytes", C.c_uint64)]


class BatchOut(C.Structure):
    _fields_ = [("out", C.c_void_p), ("out_bytes", C.c_uint64), ("consumed", C.c_uint64),
                ("n_records", C.c_uint64), ("n_selected", C.c_uint64), ("results", C.c_void_p),
                ("bad_record", C.c_int64)]


class Stats(C.Structure):
    _fields_ = [("records", C.c_uint64), ("selected", C.c_uint64), ("in_bytes", C.c_uint64),
                ("out_bytes", C.c_uint64), ("batches", C.c_uint64), ("fallback_batches", C.c_uint64),
                ("h2d_ms", C.c_double),
                ("kernel_ms", C.c_double), ("d2h_ms", C.c_double), ("wall_ms", C.c_double),
                ("device_batches", C.c_uint64 * 16)]


class BenchOut(C.Structure):
    _fields_ = [("ms_total", C.c_double), ("ms_kernel", C.c_double * 8),
                ("kernel_name", (C.c_char * 32) * 8), ("n_kernels", C.c_int32),
                ("launches_per_step", C.c_int32), ("n_records", C.c_uint64),
                ("n_selected", C.c_uint64), ("in_bytes", C.c_uint64), ("out_bytes", C.c_uint64),
                ("out_checksum", C.c_uint64)]


EXPORTS = ["bstr_create", "bstr_destroy", "bstr_last_error", "bstr_run_files", "bstr_scan_batch",
           "bstr_match_reads", "bstr_bench_resident", "bstr_host_alloc", "bstr_host_free",
           "bstr_get_stats", "bstr_reset_stats", "bstr_device_count", "bstr_lanes",
           "bstr_submit_batch", "bstr_collect_batch", "bstr_gunzip", "bstr_debug_reads",
           "bstr_debug_codes_cap", "bstr_geometry_probe", "bstr_pcie_probe", "bstr_scan_batch_noalloc", "bstr_gunzip_gpu",
           "bstr_kmers_create", "bstr_kmers_add_fastq", "bstr_kmers_add_resident", "bstr_kmers_finish", "bstr_kmers_stats", "bstr_kmers_destroy", "bstr_kmers_extend", "bstr_merge_reads"]

_lib = None


class BstrError(RuntimeError):
    def __init__(self, code, message, bad_record=-1):
        super().__init__("libbaitstr_select error %d: %s" % (code, message))
        self.code, self.bad_record = code, bad_record


def load_library():
    """Load libbaitstr_select.so (built in-tree by baitstr_b200/csrc/Makefile)."""
    global _lib
    if _lib is not None:
        return _lib
    path = library_path()
    if not os.path.exists(path):
        raise ImportError("%s is missing: build it with `make -C baitstr_b200/csrc` "
                          "(there is no CPU fallback)" % path)
    lib = C.CDLL(path)
    lib.bstr_create.argtypes = [C.POINTER(Params), C.c_char_p, C.c_size_t]
    lib.bstr_create.restype = C.c_void_p
    lib.bstr_destroy.argtypes = [C.c_void_p]
    lib.bstr_destroy.restype = None
    lib.bstr_last_error.argtypes = [C.c_void_p]
    lib.bstr_last_error.restype = C.c_char_p
    lib.bstr_run_files.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.c_int, C.c_int, C.c_int]
    lib.bstr_run_files.restype = C.c_int
    lib.bstr_scan_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.POINTER(BatchOut)]
    lib.bstr_scan_batch.restype = C.c_int
    lib.bstr_scan_batch_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p, C.c_uint64,
                                            C.POINTER(BatchOut)]
    lib.bstr_scan_batch_noalloc.restype = C.c_int
    lib.bstr_lanes.argtypes = [C.c_void_p]
    lib.bstr_lanes.restype = C.c_int
    lib.bstr_submit_batch.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_uint64, C.c_int]
    lib.bstr_submit_batch.restype = C.c_int
    lib.bstr_collect_batch.argtypes = [C.c_void_p, C.c_int, C.POINTER(BatchOut)]
    lib.bstr_collect_batch.restype = C.c_int
    lib.bstr_match_reads.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
    lib.bstr_match_reads.restype = C.c_int
    lib.bstr_debug_reads.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p,
                                     C.c_void_p]
    lib.bstr_debug_reads.restype = C.c_int
    lib.bstr_debug_codes_cap.argtypes = [C.c_void_p, C.c_uint32]
    lib.bstr_debug_codes_cap.restype = C.c_uint32
    lib.bstr_bench_resident.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64),
                                        C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(BenchOut)]
    lib.bstr_bench_resident.restype = C.c_int
    lib.bstr_pcie_probe.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64),
                                    C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double)]
    lib.bstr_pcie_probe.restype = C.c_int
    lib.bstr_host_alloc.argtypes = [C.c_size_t]
    lib.bstr_host_alloc.restype = C.c_void_p
    lib.bstr_host_free.argtypes = [C.c_void_p]
    lib.bstr_host_free.restype = None
    lib.bstr_get_stats.argtypes = [C.c_void_p, C.POINTER(Stats)]
    lib.bstr_get_stats.restype = None
    lib.bstr_reset_stats.argtypes = [C.c_void_p]
    lib.bstr_reset_stats.restype = None
    lib.bstr_device_count.argtypes = []
    lib.bstr_device_count.restype = C.c_int
    lib.bstr_gunzip.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64), C.c_char_p, C.c_size_t]
    lib.bstr_gunzip.restype = C.c_int
    lib.bstr_merge_reads.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_uint64)]
    lib.bstr_merge_reads.restype = C.c_int
    lib.bstr_kmers_create.argtypes = [C.c_void_p, C.c_int, C.c_uint64]
    lib.bstr_kmers_create.restype = C.c_void_p
    lib.bstr_kmers_add_fastq.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
    lib.bstr_kmers_add_fastq.restype = C.c_int
    lib.bstr_kmers_add_resident.argtypes = [C.c_void_p, C.c_uint64]
    lib.bstr_kmers_add_resident.restype = C.c_int
    lib.bstr_kmers_finish.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
    lib.bstr_kmers_finish.restype = C.c_int
    lib.bstr_kmers_stats.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
    lib.bstr_kmers_stats.restype = None
    lib.bstr_kmers_extend.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.c_int, C.POINTER(C.c_uint64)]
    lib.bstr_kmers_extend.restype = C.c_int
    lib.bstr_kmers_destroy.argtypes = [C.c_void_p]
    lib.bstr_kmers_destroy.restype = None
    lib.bstr_gunzip_gpu.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64), C.c_char_p, C.c_size_t]
    lib.bstr_gunzip_gpu.restype = C.c_int
    _lib = lib
    return lib


def device_count():
    return load_library().bstr_device_count()


def gunzip(path, out_fd=-1, threads=4, range_bytes=1 << 20, seg_bytes=8 << 20):
    """bstr_gunzip: inflate `path` with the parallel inflater of the ingest side (host only).
    Returns the stats dict; raises BstrError (not gzip / corrupt / CRC mismatch / truncated)."""
    lib = load_library()
    st = (C.c_uint64 * 6)()
    err = C.create_string_buffer(512)
    rc = lib.bstr_gunzip(os.fsencode(path), out_fd, threads, range_bytes, seg_bytes, st, err, len(err))
    if rc != 0:
        raise BstrError(rc, err.value.decode("latin-1"))
    return dict(zip(("in_bytes", "out_bytes", "ranges", "accepted", "redone", "members"), [int(x) for x in st]))


def gunzip_gpu(path, out_fd=-1, device=0, window_bytes=0, range_bytes=0):
    """bstr_gunzip_gpu: inflate `path` with the GPU inflater (csrc/ginflate.cu) on CUDA device `device`."""
    lib = load_library()
    st = (C.c_uint64 * 6)()
    err = C.create_string_buffer(512)
    rc = lib.bstr_gunzip_gpu(os.fsencode(path), out_fd, device, window_bytes, range_bytes, st, err, len(err))
    if rc != 0:
        raise BstrError(rc, err.value.decode("latin-1"))
    return dict(zip(("in_bytes", "out_bytes", "ranges", "accepted", "redone", "members"), [int(x) for x in st]))


def kmask_of(ks):
    m = 0
    for k in (ks or [2, 3, 4, 5, 6]):
        k = int(k)
        if 2 <= k <= 6:
            m |= 1 << k
    return m


class PinnedBuffer:
    """Page-locked host memory from bstr_host_alloc, exposed as a numpy uint8 array."""

    def __init__(self, nbytes):
        self._lib = load_library()
        self.nbytes = int(nbytes)
        self.ptr = self._lib.bstr_host_alloc(self.nbytes)
        if not self.ptr:
            raise MemoryError("bstr_host_alloc(%d) failed" % nbytes)
        self.array = np.ctypeslib.as_array((C.c_uint8 * self.nbytes).from_address(self.ptr))

    def close(self):
        if self.ptr:
            self.array = None
            self._lib.bstr_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class KmerCounter:
    """bstr_kmers_*: the k-mer table of extend_STR_reads on the GPU (include/baitstr_select.h)."""

    def __init__(self, ctx, klen, expected_kmers=1 << 20):
        self._ctx, self._lib = ctx, ctx._lib
        self._h = self._lib.bstr_kmers_create(ctx._ctx, int(klen), int(expected_kmers))
        if not self._h:
            ctx._raise(-3)

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def close(self):
        if self._h:
            self._lib.bstr_kmers_destroy(self._h)
            self._h = None

    def _check(self, rc):
        if rc:
            self._ctx._raise(rc)

    def add_fastq(self, raw):
        addr, n, keep = Context._addr(raw)
        self._check(self._lib.bstr_kmers_add_fastq(self._h, addr, n))

    def add_resident(self, nbytes):
        """count the batch bstr_scan_batch has just scanned on this context, where it lies in device memory"""
        self._check(self._lib.bstr_kmers_add_resident(self._h, int(nbytes)))

    def finish(self, min_threshold=2, max_threshold=255):
        kp, cp, n = C.c_void_p(), C.c_void_p(), C.c_uint64()
        self._check(self._lib.bstr_kmers_finish(self._h, int(min_threshold), int(max_threshold), C.byref(kp), C.byref(cp), C.byref(n)))
        if not n.value:
            return np.zeros(0, np.uint64), np.zeros(0, np.uint8)
        return (np.ctypeslib.as_array((C.c_uint64 * n.value).from_address(kp.value)).copy(),
                np.ctypeslib.as_array((C.c_uint8 * n.value).from_address(cp.value)).copy())

    def extend(self, blocks, min_threshold=2, max_threshold=255, flanks=1024):
        import tempfile
        addr, n, keep = Context._addr(blocks)
        st = (C.c_uint64 * 4)()
        with tempfile.TemporaryFile() as f:
            self._check(self._lib.bstr_kmers_extend(self._h, int(min_threshold), int(max_threshold), int(flanks), addr, n, f.fileno(), st))
            f.seek(0)
            out = f.read()
        return out, dict(zip(("blocks", "contigs", "ends_not_extended", "kmers_kept"), [int(x) for x in st]))

    def stats(self):
        st = (C.c_uint64 * 4)()
        self._lib.bstr_kmers_stats(self._h, st)
        return dict(zip(("records", "kmers", "distinct", "slots"), [int(x) for x in st]))


class Context:
    """One scan context = the reference's Pool + compiled patterns
    (/root/reference/src/select_STR_reads.pyx:347-357) for one flag set."""

    def __init__(self, n=2, f=31, ks=None, rmhm=False, altalgo=True, illumina=False,
                 devices=None, want_results=False, batch_bytes=0, host_threads=0):
        self._lib = load_library()
        p = Params()
        p.abi_version = ABI_VERSION
        p.num_copies, p.flank, p.kmask = int(n), int(f), kmask_of(ks)
        p.remove_homopolymers, p.altalgo, p.illumina_quals = int(bool(rmhm)), int(bool(altalgo)), int(bool(illumina))
        devices = list(devices or [])
        p.n_devices = len(devices)
        for i, d in enumerate(devices[:16]):
            p.devices[i] = int(d)
        p.host_threads, p.want_results, p.batch_bytes = int(host_threads), int(bool(want_results)), int(batch_bytes)
        err = C.create_string_buffer(512)
        self._ctx = self._lib.bstr_create(C.byref(p), err, len(err))
        if not self._ctx:
            raise BstrError(-1, err.value.decode("latin-1", "replace"))
        self.want_results = bool(want_results)

    def close(self):
        if getattr(self, "_ctx", None):
            self._lib.bstr_destroy(self._ctx)
            self._ctx = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _raise(self, rc, bad=-1):
        raise BstrError(rc, self._lib.bstr_last_error(self._ctx).decode("latin-1", "replace"), bad)

    @staticmethod
    def _addr(raw):
        if isinstance(raw, PinnedBuffer):
            return raw.ptr, raw.nbytes, raw
        if isinstance(raw, np.ndarray):
            a = np.ascontiguousarray(raw, dtype=np.uint8)
            return a.ctypes.data, a.nbytes, a
        b = bytes(raw)
        a = np.frombuffer(b, dtype=np.uint8)
        return (a.ctypes.data if len(b) else None), len(b), a

    def scan_batch(self, raw, final=True, nbytes=None):
        """raw FASTQ bytes -> (output bytes, info dict).  H2D, kernels and D2H inside."""
        addr, n, keep = self._addr(raw)
        if nbytes is not None:
            n = int(nbytes)
        bo = BatchOut()
        rc = self._lib.bstr_scan_batch(self._ctx, addr, n, int(bool(final)), C.byref(bo))
        if rc:
            self._raise(rc, bo.bad_record)
        out = C.string_at(bo.out, bo.out_bytes) if bo.out_bytes else b""
        info = {"consumed": bo.consumed, "n_records": bo.n_records, "n_selected": bo.n_selected,
                "out_bytes": bo.out_bytes}
        if self.want_results:
            if bo.n_records:
                arr = np.ctypeslib.as_array((C.c_uint32 * (4 * bo.n_records)).from_address(bo.results))
                info["results"] = arr.reshape(-1, 4).copy()
            else:
                info["results"] = np.zeros((0, 4), dtype=np.uint32)
        return out, info

    def scan_batch_noalloc(self, raw, final=True, out_capacity=None):
        """bstr_scan_batch_noalloc: as scan_batch, but the library allocates nothing (static device
        scratch; raw at most NOALLOC_MAX_BYTES).  The output lands in a caller-owned numpy buffer."""
        addr, n, keep = self._addr(raw)
        cap = int(2 * n + 4096 if out_capacity is None else out_capacity)
        out = np.empty(max(cap, 1), dtype=np.uint8)
        bo = BatchOut()
        rc = self._lib.bstr_scan_batch_noalloc(self._ctx, addr, n, int(bool(final)), out.ctypes.data, cap, C.byref(bo))
        info = {"consumed": bo.consumed, "n_records": bo.n_records, "n_selected": bo.n_selected,
                "out_bytes": bo.out_bytes}
        if rc:
            err = BstrError(rc, self._lib.bstr_last_error(self._ctx).decode("latin-1", "replace"), bo.bad_record)
            err.info = info
            raise err
        return out[:bo.out_bytes].tobytes(), info

    def scan_batch_nocopy(self, raw, final=True, nbytes=None):
        """Like scan_batch but leaves the output in the library's pinned buffer (bench)."""
        addr, n, keep = self._addr(raw)
        if nbytes is not None:
            n = int(nbytes)
        bo = BatchOut()
        rc = self._lib.bstr_scan_batch(self._ctx, addr, n, int(bool(final)), C.byref(bo))
        if rc:
            self._raise(rc, bo.bad_record)
        return bo

    def lanes(self):
        n = self._lib.bstr_lanes(self._ctx)
        if n < 0:
            self._raise(n)
        return n

    def submit(self, lane, raw, final=True, nbytes=None, offset=0):
        """Enqueue one batch on `lane` (no wait).  `raw` must stay alive until collect()."""
        addr, n, keep = self._addr(raw)
        if nbytes is not None:
            n = int(nbytes)
        rc = self._lib.bstr_submit_batch(self._ctx, int(lane), (addr or 0) + int(offset), n, int(bool(final)))
        if rc:
            self._raise(rc)
        return keep

    def collect(self, lane, copy=True):
        """Wait for the batch on `lane`; returns (output bytes or BatchOut, info)."""
        bo = BatchOut()
        rc = self._lib.bstr_collect_batch(self._ctx, int(lane), C.byref(bo))
        if rc:
            self._raise(rc, bo.bad_record)
        info = {"consumed": bo.consumed, "n_records": bo.n_records, "n_selected": bo.n_selected,
                "out_bytes": bo.out_bytes}
        if not copy:
            return bo, info
        return (C.string_at(bo.out, bo.out_bytes) if bo.out_bytes else b""), info

    def match_reads(self, seqs):
        """list of bytes -> uint32 [n, 4] (fstart, rstart, fck, rck); fck = copies << 3 | k."""
        nreads = len(seqs)
        off = np.zeros(nreads + 1, dtype=np.uint64)
        if nreads:
            off[1:] = np.cumsum([len(s) for s in seqs])
        blob = np.frombuffer(b"".join(seqs) + b"\0", dtype=np.uint8)
        res = np.zeros((nreads, 4), dtype=np.uint32)
        rc = self._lib.bstr_match_reads(self._ctx, blob.ctypes.data, off.ctypes.data, nreads,
                                        res.ctypes.data)
        if rc:
            self._raise(rc)
        return res

    def debug_reads(self, seqs):
        """list of bytes -> per read (info[8], forward codes, reverse-strand codes): the -d report
        of include/baitstr_select.h: bstr_debug_reads."""
        nreads = len(seqs)
        off = np.zeros(nreads + 1, dtype=np.uint64)
        coff = np.zeros(nreads + 1, dtype=np.uint64)
        if nreads:
            off[1:] = np.cumsum([len(s) for s in seqs])
            caps = [2 * self._lib.bstr_debug_codes_cap(self._ctx, len(s)) for s in seqs]
            coff[1:] = np.cumsum(caps)
        blob = np.frombuffer(b"".join(seqs) + b"\0", dtype=np.uint8)
        codes = np.zeros(int(coff[-1]) + 1, dtype=np.uint8)
        info = np.zeros((nreads, 8), dtype=np.uint32)
        rc = self._lib.bstr_debug_reads(self._ctx, blob.ctypes.data, off.ctypes.data, nreads,
                                        coff.ctypes.data, codes.ctypes.data, info.ctypes.data)
        if rc:
            self._raise(rc)
        out = []
        for i in range(nreads):
            lo, hi = int(coff[i]), int(coff[i + 1])
            half = (hi - lo) // 2
            out.append((info[i], codes[lo:lo + int(info[i, 0])], codes[lo + half:lo + half + int(info[i, 4])]))
        return out

    def run_files(self, paths, out_fd=1, err_fd=2):
        arr = (C.c_char_p * len(paths))(*[os.fsencode(p) for p in paths])
        rc = self._lib.bstr_run_files(self._ctx, arr, len(paths), int(out_fd), int(err_fd))
        if rc:
            self._raise(rc)

    def bench_resident(self, batches, steps, warmup, flush_l2=False):
        """batches: list of raw FASTQ buffers (each whole records).  One step = the device
        pipeline over every batch, inputs resident in HBM."""
        keep = [self._addr(b) for b in batches]
        ptrs = (C.c_void_p * len(keep))(*[k[0] for k in keep])
        sizes = (C.c_uint64 * len(keep))(*[k[1] for k in keep])
        bo = BenchOut()
        rc = self._lib.bstr_bench_resident(self._ctx, ptrs, sizes, len(keep), int(steps),
                                           int(warmup), int(bool(flush_l2)), C.byref(bo))
        if rc:
            self._raise(rc)
        kern = {bo.kernel_name[i].value.decode(): bo.ms_kernel[i] for i in range(bo.n_kernels)}
        return {"ms_total": bo.ms_total, "kernels_ms": kern, "launches_per_step": bo.launches_per_step,
                "n_records": bo.n_records, "n_selected": bo.n_selected, "in_bytes": bo.in_bytes,
                "out_bytes": bo.out_bytes, "out_checksum": bo.out_checksum}

    def pcie_probe(self, jobs, out_bytes, mode, reps=1):
        """bstr_pcie_probe: jobs = [(buffer, offset, nbytes)], out_bytes per job; mode 1 = H2D,
        2 = D2H, 3 = both.  Returns the wall time in ms."""
        keep = [self._addr(b) for b, _, _ in jobs]
        ptrs = (C.c_void_p * len(jobs))(*[(k[0] or 0) + int(j[1]) for k, j in zip(keep, jobs)])
        sizes = (C.c_uint64 * len(jobs))(*[int(j[2]) for j in jobs])
        outs = (C.c_uint64 * len(jobs))(*[int(x) for x in out_bytes])
        ms = C.c_double(0)
        rc = self._lib.bstr_pcie_probe(self._ctx, ptrs, sizes, outs, len(jobs), int(mode), int(reps), C.byref(ms))
        if rc:
            self._raise(rc)
        return ms.value

    def merge_reads(self, fastq, klength, min_threshold=4, max_threshold=10000, include_all=False):
        """bstr_merge_reads: the select stage's output -> (the blocks merge_STR_reads would print, stats dict)."""
        import tempfile
        addr, n, keep = self._addr(fastq)
        st = (C.c_uint64 * 6)()
        with tempfile.TemporaryFile() as f:
            rc = self._lib.bstr_merge_reads(self._ctx, addr, n, int(klength), int(min_threshold), int(max_threshold),
                                            int(bool(include_all)), f.fileno(), st)
            if rc:
                self._raise(rc)
            f.seek(0)
            out = f.read()
        return out, dict(zip(("reads", "keys", "chains", "blocks", "printed", "longest_chain"), [int(x) for x in st]))

    def kmer_counter(self, klen, expected_kmers=1 << 20):
        """a bstr_kmers handle on this context (use as a context manager)"""
        return KmerCounter(self, klen, expected_kmers)

    def extend_reads(self, blocks, read_batches, klen, min_threshold=2, max_threshold=255, flanks=1024, expected_kmers=1 << 20):
        """bstr_kmers_* + bstr_kmers_extend: the extend stage.  blocks = the merge stage's output, read_batches =
        raw FASTQ batches of the sample's reads (whole records each) -> (the contigs extend_STR_reads would print, stats)."""
        if klen % 2 == 0:
            klen -= 1                               # extend_STR_reads.c:1072-1075
        with self.kmer_counter(klen, expected_kmers) as kc:
            for b in [blocks] + list(read_batches):   # the reference counts its argv[4 ..]: the blocks' file first
                kc.add_fastq(b)
            return kc.extend(blocks, min_threshold, max_threshold, flanks)

    def count_kmers(self, batches, klen, min_threshold=2, max_threshold=255, expected_kmers=1 << 20):
        """bstr_kmers_*: the k-mer counting pass of extend_STR_reads over raw FASTQ batches (whole
        records each).  Returns (uint64 k-mers sorted, uint8 counts, stats dict)."""
        with self.kmer_counter(klen, expected_kmers) as kc:
            for b in batches:
                kc.add_fastq(b)
            kmers, counts = kc.finish(min_threshold, max_threshold)
            return kmers, counts, kc.stats()

    def stats(self):
        st = Stats()
        self._lib.bstr_get_stats(self._ctx, C.byref(st))
        d = {k: getattr(st, k) for k, _ in Stats._fields_}
        d["device_batches"] = [int(x) for x in st.device_batches]
        return d

    def reset_stats(self):
        self._lib.bstr_reset_stats(self._ctx)
