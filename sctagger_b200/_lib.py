"""ctypes binding of libsctagger_b200.so (C ABI: include/sctagger_b200.h).

This is the only module that touches the shared library.  There is no CPU fallback: if the library
is missing, or no CUDA device is visible, every compute entry point raises.
"""
from __future__ import annotations

import ctypes
import os
import threading

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("SCT_LIBRARY") or os.path.join(_HERE, "libsctagger_b200.so")   # SCT_LIBRARY: a build variant (tuning)

# every symbol include/sctagger_b200.h declares (checked by tests/test_abi.py)
EXPORTS = (
    "sct_device_count", "sct_last_error", "sct_create", "sct_destroy", "sct_match", "sct_stage",
    "sct_run_staged", "sct_fetch_staged", "sct_stats", "sct_set_option", "sct_chunk_capacity",
    "sct_measure_int32_peak", "sct_flush_l2",
    "sct_host_alloc", "sct_host_free", "sct_host_register", "sct_host_unregister", "sct_work_counters", "sct_kernel_name",
    "sct_align_create", "sct_align_destroy", "sct_align", "sct_align_stage", "sct_align_run_staged",
    "sct_align_stats", "sct_align_set_option", "sct_align_strand_dist_noalloc",
    "sct_io_read_long_reads", "sct_io_read_fastqs", "sct_io_reads_info", "sct_io_reads_free",
    "sct_io_format_matches", "sct_io_format_lr_bc", "sct_io_buffer_info", "sct_io_buffer_free",
    "sct_io_read_lr_bc_segments", "sct_io_read_whitelist",
    "sct_wl_create", "sct_wl_destroy", "sct_wl_count", "sct_wl_result", "sct_wl_reset", "sct_wl_stage",
    "sct_wl_run_staged", "sct_wl_stats",
)

_u8p = ctypes.POINTER(ctypes.c_uint8)
_i32p = ctypes.POINTER(ctypes.c_int32)
_i64p = ctypes.POINTER(ctypes.c_int64)
_f32p = ctypes.POINTER(ctypes.c_float)
_vp = ctypes.c_void_p

_lib = None
_lock = threading.Lock()


class SctError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"sctagger_b200 error {code}: {msg}")
        self.code = code


def lib():
    """Load the shared library (once).  Raises OSError with a build hint if it is missing."""
    global _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(SO_PATH):
                raise OSError(f"{SO_PATH} not found: build it with sctagger_b200/csrc/build.sh "
                              f"(or __graft_entry__.build()); there is no CPU fallback")
            L = ctypes.CDLL(SO_PATH)
            L.sct_device_count.restype = ctypes.c_int
            L.sct_device_count.argtypes = []
            L.sct_last_error.restype = ctypes.c_char_p
            L.sct_last_error.argtypes = []
            L.sct_create.restype = ctypes.c_int
            L.sct_create.argtypes = [ctypes.c_int, _u8p, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                     ctypes.c_int32, ctypes.c_double, ctypes.POINTER(_vp)]
            L.sct_destroy.restype = ctypes.c_int
            L.sct_destroy.argtypes = [_vp]
            L.sct_match.restype = ctypes.c_int
            L.sct_match.argtypes = [_vp, _u8p, _i64p, ctypes.c_int64, _i32p, _i32p, _i64p,
                                    ctypes.POINTER(_i32p), _i64p]
            L.sct_stage.restype = ctypes.c_int
            L.sct_stage.argtypes = [_vp, _u8p, _i64p, ctypes.c_int64]
            L.sct_run_staged.restype = ctypes.c_int
            L.sct_run_staged.argtypes = [_vp, _f32p, _f32p]
            L.sct_fetch_staged.restype = ctypes.c_int
            L.sct_fetch_staged.argtypes = [_vp, _i32p, _i32p, _i64p, ctypes.POINTER(_i32p), _i64p]
            L.sct_stats.restype = ctypes.c_int
            L.sct_stats.argtypes = [_vp, _i64p]
            L.sct_set_option.restype = ctypes.c_int
            L.sct_set_option.argtypes = [_vp, ctypes.c_char_p, ctypes.c_int64]
            L.sct_chunk_capacity.restype = ctypes.c_int
            L.sct_chunk_capacity.argtypes = [_vp, _i64p, _i64p]
            L.sct_measure_int32_peak.restype = ctypes.c_int
            L.sct_measure_int32_peak.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                                 ctypes.POINTER(ctypes.c_double)]
            L.sct_flush_l2.restype = ctypes.c_int
            L.sct_flush_l2.argtypes = [_vp]
            L.sct_host_alloc.restype = ctypes.c_int
            L.sct_host_alloc.argtypes = [ctypes.c_size_t, ctypes.POINTER(_vp)]
            L.sct_host_free.restype = ctypes.c_int
            L.sct_host_free.argtypes = [_vp]
            L.sct_host_register.restype = ctypes.c_int
            L.sct_host_register.argtypes = [_vp, ctypes.c_size_t]
            L.sct_host_unregister.restype = ctypes.c_int
            L.sct_host_unregister.argtypes = [_vp]
            L.sct_work_counters.restype = ctypes.c_int
            L.sct_work_counters.argtypes = [_vp, _i64p]
            L.sct_kernel_name.restype = ctypes.c_char_p
            L.sct_kernel_name.argtypes = [_vp]
            _i8p = ctypes.POINTER(ctypes.c_int8)
            L.sct_align_create.restype = ctypes.c_int
            L.sct_align_create.argtypes = [ctypes.c_int, _u8p, ctypes.c_int32, ctypes.c_int32, ctypes.c_double,
                                           ctypes.POINTER(_vp)]
            L.sct_align_destroy.restype = ctypes.c_int
            L.sct_align_destroy.argtypes = [_vp]
            L.sct_align.restype = ctypes.c_int
            L.sct_align.argtypes = [_vp, _u8p, _i64p, ctypes.c_int64, _i8p, _i32p, _i64p, ctypes.POINTER(_i32p), _i64p]
            L.sct_align_stage.restype = ctypes.c_int
            L.sct_align_stage.argtypes = [_vp, _u8p, _i64p, ctypes.c_int64]
            L.sct_align_run_staged.restype = ctypes.c_int
            L.sct_align_run_staged.argtypes = [_vp, _f32p, _f32p]
            L.sct_align_stats.restype = ctypes.c_int
            L.sct_align_stats.argtypes = [_vp, _i64p]
            L.sct_align_set_option.restype = ctypes.c_int
            L.sct_align_set_option.argtypes = [_vp, ctypes.c_char_p, ctypes.c_int64]
            L.sct_io_read_long_reads.restype = ctypes.c_int
            L.sct_io_read_long_reads.argtypes = [ctypes.c_char_p, ctypes.c_int32, ctypes.POINTER(_vp)]
            for f in (L.sct_io_read_lr_bc_segments, L.sct_io_read_whitelist):
                f.restype = ctypes.c_int
                f.argtypes = [ctypes.c_char_p, ctypes.c_int32, ctypes.POINTER(_vp)]
            L.sct_wl_create.restype = ctypes.c_int
            L.sct_wl_create.argtypes = [ctypes.c_int, _u8p, _i64p, ctypes.c_int64, ctypes.c_int32, ctypes.POINTER(_vp)]
            L.sct_wl_destroy.restype = ctypes.c_int
            L.sct_wl_destroy.argtypes = [_vp]
            L.sct_wl_count.restype = ctypes.c_int
            L.sct_wl_count.argtypes = [_vp, _u8p, _i64p, ctypes.c_int64, _i32p, _i32p]
            L.sct_wl_result.restype = ctypes.c_int
            L.sct_wl_result.argtypes = [_vp, _i64p, ctypes.POINTER(_i32p), ctypes.POINTER(_i64p)]
            L.sct_wl_reset.restype = ctypes.c_int
            L.sct_wl_reset.argtypes = [_vp]
            L.sct_wl_stage.restype = ctypes.c_int
            L.sct_wl_stage.argtypes = [_vp, _u8p, _i64p, ctypes.c_int64]
            L.sct_wl_run_staged.restype = ctypes.c_int
            L.sct_wl_run_staged.argtypes = [_vp, _f32p]
            L.sct_wl_stats.restype = ctypes.c_int
            L.sct_wl_stats.argtypes = [_vp, _i64p]
            L.sct_io_read_fastqs.restype = ctypes.c_int
            L.sct_io_read_fastqs.argtypes = [ctypes.c_char_p, ctypes.c_int32, ctypes.c_int32, ctypes.POINTER(_vp)]
            L.sct_io_reads_info.restype = ctypes.c_int
            L.sct_io_reads_info.argtypes = [_vp, _i64p, ctypes.POINTER(_u8p), ctypes.POINTER(_i64p),
                                            ctypes.POINTER(_u8p), ctypes.POINTER(_i64p)]
            L.sct_io_reads_free.restype = ctypes.c_int
            L.sct_io_reads_free.argtypes = [_vp]
            L.sct_io_format_matches.restype = ctypes.c_int
            L.sct_io_format_matches.argtypes = [_u8p, _i64p, _u8p, _i64p, ctypes.c_int64, _i32p, _i32p, _i64p, _i32p,
                                                _u8p, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                                ctypes.POINTER(_vp)]
            L.sct_io_format_lr_bc.restype = ctypes.c_int
            L.sct_io_format_lr_bc.argtypes = [_u8p, _i64p, _u8p, _i64p, ctypes.c_int64, _i64p, _i64p, _i64p, _i64p,
                                              ctypes.c_int32, ctypes.c_int32, ctypes.POINTER(_vp)]
            L.sct_io_buffer_info.restype = ctypes.c_int
            L.sct_io_buffer_info.argtypes = [_vp, ctypes.POINTER(_u8p), _i64p, _i64p]
            L.sct_io_buffer_free.restype = ctypes.c_int
            L.sct_io_buffer_free.argtypes = [_vp]
            _lib = L
    return _lib


E_IO, E_FORMAT = -6, -7


def _check(rc):
    if rc != 0:
        msg = lib().sct_last_error().decode("utf-8", "replace")
        if rc == E_FORMAT:          # the reference's IndexError on a short row (scTagger.py:613-614)
            raise IndexError(msg)
        if rc == E_IO:              # the reference's open()/gzip.open() failures
            raise OSError(msg)
        raise SctError(rc, msg)


def device_count() -> int:
    return int(lib().sct_device_count())


def _ptr(a, ct):
    return a.ctypes.data_as(ctypes.POINTER(ct))


def _as_u8(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    return a if a.size else np.zeros(1, dtype=np.uint8)


STAT_NAMES = ("launches", "fast_segments", "generic_segments", "h2d_bytes", "d2h_bytes", "chunks",
              "dp_segments", "seed_candidates")
WORK_NAMES = ("probes", "verifications", "prechecks", "phases_run", "wide_segments", "automaton_lop3_lane_ops", "reserved6",
              "reserved7")


class Pinned:
    """One page-locked host allocation (sct_host_alloc) exposed as a numpy array (`.array`).  Arrays living in such
    memory are copied from / to by the GPU's copy engines directly: sct_match skips its staging pass for them.  The
    memory is released by close() or when the object is collected — keep the object alive as long as the array is used."""

    def __init__(self, shape, dtype):
        dtype = np.dtype(dtype)
        shape = (int(shape),) if np.isscalar(shape) else tuple(int(x) for x in shape)
        count = int(np.prod(shape)) if shape else 1
        self.nbytes = max(64, count * dtype.itemsize)
        self._p = _vp()
        _check(lib().sct_host_alloc(self.nbytes, ctypes.byref(self._p)))
        buf = (ctypes.c_uint8 * self.nbytes).from_address(self._p.value)
        self.array = np.frombuffer(buf, dtype=dtype, count=count).reshape(shape)

    @classmethod
    def copy_of(cls, a):
        a = np.asarray(a)
        p = cls(a.shape, a.dtype)
        np.copyto(p.array, a)
        return p

    def close(self):
        if getattr(self, "_p", None) is not None and self._p.value:
            self.array = None
            lib().sct_host_free(self._p)
            self._p = _vp()

    __del__ = close


class ResultBuffers:
    """Caller-owned, reusable, page-locked result arrays for Matcher.match(..., out=...): the device writes min_ed, n_ties
    and tie_offsets straight into them (no staging, no host pass).  Capacity in segments; reuse it across calls — the
    arrays of one call are overwritten by the next call that uses the same buffers."""

    def __init__(self, capacity: int):
        self.capacity = int(capacity)
        self._min = Pinned(max(self.capacity, 1), np.int32)
        self._nt = Pinned(max(self.capacity, 1), np.int32)
        self._toff = Pinned(self.capacity + 1, np.int64)

    def views(self, n):
        if n > self.capacity:
            raise ValueError(f"ResultBuffers of capacity {self.capacity} cannot hold {n} segments")
        return self._min.array[:max(n, 1)], self._nt.array[:max(n, 1)], self._toff.array[:n + 1]

    def close(self):
        for x in (self._min, self._nt, self._toff):
            x.close()


class AlignBuffers:
    """Caller-owned, reusable, page-locked result arrays for Aligner.align(..., out=...): strand, dist and loc_offsets are
    written by the device directly.  Capacity in reads."""

    def __init__(self, capacity: int):
        self.capacity = int(capacity)
        self._strand = Pinned(max(self.capacity, 1), np.int8)
        self._dist = Pinned(max(self.capacity, 1), np.int32)
        self._locoff = Pinned(self.capacity + 1, np.int64)

    def views(self, n):
        if n > self.capacity:
            raise ValueError(f"AlignBuffers of capacity {self.capacity} cannot hold {n} reads")
        return self._strand.array[:max(n, 1)], self._dist.array[:max(n, 1)], self._locoff.array[:n + 1]

    def close(self):
        for x in (self._strand, self._dist, self._locoff):
            x.close()


class Matcher:
    """One handle = one GPU with the barcode table resident (sct_create .. sct_destroy)."""

    def __init__(self, barcodes: np.ndarray, max_error: int, threads: int = 16, mem_gb: float = 16.0,
                 device: int = 0, **options):
        barcodes = np.ascontiguousarray(barcodes, dtype=np.uint8)
        if barcodes.ndim != 2:
            raise ValueError("barcodes must be a (B, bl) uint8 matrix")
        self._h = _vp()
        self.n_barcodes, self.barcode_length = int(barcodes.shape[0]), int(barcodes.shape[1])
        self.max_error = int(max_error)
        self.device = int(device)
        _check(lib().sct_create(int(device), _ptr(_as_u8(barcodes), ctypes.c_uint8), self.n_barcodes,
                                self.barcode_length, int(max_error), int(threads), float(mem_gb),
                                ctypes.byref(self._h)))
        for k, v in options.items():
            self.set_option(k, v)

    def set_option(self, key: str, value: int):
        _check(lib().sct_set_option(self._h, key.encode(), int(value)))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib().sct_destroy(self._h)
            self._h = _vp()

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # -- results ------------------------------------------------------------------------------------
    @staticmethod
    def _out(n):
        return (np.empty(max(n, 1), dtype=np.int32), np.empty(max(n, 1), dtype=np.int32),
                np.empty(n + 1, dtype=np.int64))

    @staticmethod
    def _ties(ids, total):
        if total == 0:
            return np.zeros(0, dtype=np.int32)
        return np.ctypeslib.as_array(ids, shape=(total,)).copy()

    def match(self, seg_bytes: np.ndarray, offsets: np.ndarray, out: "ResultBuffers" = None, copy_ties: bool = True):
        """sct_match: host buffers in, host buffers out (chunked, double-buffered async copies inside).
        Arrays in page-locked memory (`Pinned`, `ResultBuffers`) are copied from / to directly; pageable numpy arrays
        are staged by the library's host threads.  With `out`, the result arrays are views of the caller's
        ResultBuffers; with copy_ties=False, tie_ids is a view of the library-owned buffer, valid until the next call
        on this Matcher (the C ABI's own contract)."""
        from .match_trie import MatchResult
        sb = _as_u8(seg_bytes)
        off = np.ascontiguousarray(offsets, dtype=np.int64)
        n = off.shape[0] - 1
        min_ed, n_ties, tie_off = self._out(n) if out is None else out.views(n)
        ids = _i32p()
        total = ctypes.c_int64(0)
        _check(lib().sct_match(self._h, _ptr(sb, ctypes.c_uint8), _ptr(off, ctypes.c_int64), n,
                               _ptr(min_ed, ctypes.c_int32), _ptr(n_ties, ctypes.c_int32),
                               _ptr(tie_off, ctypes.c_int64), ctypes.byref(ids), ctypes.byref(total)))
        t = int(total.value)
        if copy_ties or t == 0:
            ties = self._ties(ids, t)
        else:
            ties = np.ctypeslib.as_array(ids, shape=(t,))
        return MatchResult(min_ed[:n], n_ties[:n], tie_off, ties)

    def stage(self, seg_bytes: np.ndarray, offsets: np.ndarray):
        sb = _as_u8(seg_bytes)
        off = np.ascontiguousarray(offsets, dtype=np.int64)
        self._staged_n = off.shape[0] - 1
        _check(lib().sct_stage(self._h, _ptr(sb, ctypes.c_uint8), _ptr(off, ctypes.c_int64), self._staged_n))

    def run_staged(self):
        """-> (pass_ms, match_kernel_ms): device time of one pass over the resident batch (CUDA events)."""
        a, b = ctypes.c_float(0), ctypes.c_float(0)
        _check(lib().sct_run_staged(self._h, ctypes.byref(a), ctypes.byref(b)))
        return float(a.value), float(b.value)

    def fetch_staged(self):
        from .match_trie import MatchResult
        n = self._staged_n
        min_ed, n_ties, tie_off = self._out(n)
        ids = _i32p()
        total = ctypes.c_int64(0)
        _check(lib().sct_fetch_staged(self._h, _ptr(min_ed, ctypes.c_int32), _ptr(n_ties, ctypes.c_int32),
                                      _ptr(tie_off, ctypes.c_int64), ctypes.byref(ids), ctypes.byref(total)))
        return MatchResult(min_ed[:n], n_ties[:n], tie_off, self._ties(ids, int(total.value)))

    def stats(self) -> dict:
        s = np.zeros(8, dtype=np.int64)
        _check(lib().sct_stats(self._h, _ptr(s, ctypes.c_int64)))
        return dict(zip(STAT_NAMES, s.tolist()))

    def kernel_name(self) -> str:
        return lib().sct_kernel_name(self._h).decode()

    def work_counters(self) -> dict:
        s = np.zeros(8, dtype=np.int64)
        _check(lib().sct_work_counters(self._h, _ptr(s, ctypes.c_int64)))
        return dict(zip(WORK_NAMES, s.tolist()))

    def chunk_capacity(self):
        a, b = ctypes.c_int64(0), ctypes.c_int64(0)
        _check(lib().sct_chunk_capacity(self._h, ctypes.byref(a), ctypes.byref(b)))
        return int(a.value), int(b.value)

    def flush_l2(self):
        _check(lib().sct_flush_l2(self._h))


ALIGN_STAT_NAMES = ("launches", "chunks", "h2d_bytes", "d2h_bytes", "device_ns", "reads", "reserved6", "reserved7")


class Aligner:
    """One handle = one GPU holding the adapter's Myers tables (sct_align_create .. sct_align_destroy): the
    `get_lr_bc_alns` seam of extract_lr_bc (reference scTagger.py:307-314, 176-196)."""

    def __init__(self, adapter: bytes, threads: int = 1, mem_gb: float = 16.0, device: int = 0, **options):
        ad = np.frombuffer(bytes(adapter), dtype=np.uint8).copy()
        self._h = _vp()
        self.device = int(device)
        _check(lib().sct_align_create(int(device), _ptr(_as_u8(ad), ctypes.c_uint8), int(ad.size), int(threads),
                                      float(mem_gb), ctypes.byref(self._h)))
        for k, v in options.items():
            _check(lib().sct_align_set_option(self._h, k.encode(), int(v)))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib().sct_align_destroy(self._h)
            self._h = _vp()

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def align(self, read_bytes: np.ndarray, offsets: np.ndarray, out: "AlignBuffers" = None, copy_locs: bool = True):
        """-> (strand int8[n] ('+'/'-' as ASCII, 0 = NA), dist int32[n] (-1 = NA), loc_offsets int64[n+1], locs int32[]).
        Arrays in page-locked memory (`Pinned`, `AlignBuffers`) are copied from / to directly; pageable numpy arrays are
        staged by the library's host threads.  With copy_locs=False, locs is a view of the library-owned buffer, valid
        until the next call on this Aligner."""
        rb = _as_u8(read_bytes)
        off = np.ascontiguousarray(offsets, dtype=np.int64)
        n = off.shape[0] - 1
        if out is None:
            strand = np.empty(max(n, 1), dtype=np.int8)
            dist = np.empty(max(n, 1), dtype=np.int32)
            loc_off = np.empty(n + 1, dtype=np.int64)
        else:
            strand, dist, loc_off = out.views(n)
        ids = _i32p()
        total = ctypes.c_int64(0)
        _check(lib().sct_align(self._h, _ptr(rb, ctypes.c_uint8), _ptr(off, ctypes.c_int64), n,
                               _ptr(strand, ctypes.c_int8), _ptr(dist, ctypes.c_int32), _ptr(loc_off, ctypes.c_int64),
                               ctypes.byref(ids), ctypes.byref(total)))
        t = int(total.value)
        if t == 0:
            locs = np.zeros(0, dtype=np.int32)
        elif copy_locs:
            locs = np.ctypeslib.as_array(ids, shape=(t,)).copy()
        else:
            locs = np.ctypeslib.as_array(ids, shape=(t,))
        return strand[:n], dist[:n], loc_off, locs

    def stage(self, read_bytes: np.ndarray, offsets: np.ndarray):
        rb = _as_u8(read_bytes)
        off = np.ascontiguousarray(offsets, dtype=np.int64)
        _check(lib().sct_align_stage(self._h, _ptr(rb, ctypes.c_uint8), _ptr(off, ctypes.c_int64), off.shape[0] - 1))

    def run_staged(self):
        a, b = ctypes.c_float(0), ctypes.c_float(0)
        _check(lib().sct_align_run_staged(self._h, ctypes.byref(a), ctypes.byref(b)))
        return float(a.value), float(b.value)

    def stats(self) -> dict:
        s = np.zeros(8, dtype=np.int64)
        _check(lib().sct_align_stats(self._h, _ptr(s, ctypes.c_int64)))
        return dict(zip(ALIGN_STAT_NAMES, s.tolist()))


def align_strand_dist_noalloc(adapter: bytes, read_bytes: np.ndarray, offsets: np.ndarray, device: int = 0):
    """sct_align_strand_dist_noalloc: strand + distance per read without any device allocation (fixed static scratch:
    <= 8192 reads, <= 2 MiB of bases, reads <= 16384 bases).  -> (strand int8[n], dist int32[n])"""
    L = lib()
    L.sct_align_strand_dist_noalloc.restype = ctypes.c_int
    L.sct_align_strand_dist_noalloc.argtypes = [ctypes.c_int, _u8p, ctypes.c_int32, _u8p, _i64p, ctypes.c_int64,
                                                ctypes.POINTER(ctypes.c_int8), _i32p]
    ad = np.frombuffer(bytes(adapter), dtype=np.uint8).copy()
    rb = _as_u8(read_bytes)
    off = np.ascontiguousarray(offsets, dtype=np.int64)
    n = off.shape[0] - 1
    strand = np.zeros(max(n, 1), dtype=np.int8)
    dist = np.zeros(max(n, 1), dtype=np.int32)
    _check(L.sct_align_strand_dist_noalloc(int(device), _ptr(_as_u8(ad), ctypes.c_uint8), int(ad.size), _ptr(rb, ctypes.c_uint8),
                                           _ptr(off, ctypes.c_int64), n, _ptr(strand, ctypes.c_int8),
                                           _ptr(dist, ctypes.c_int32)))
    return strand[:n], dist[:n]


def align_reads(adapter: bytes, read_bytes: np.ndarray, offsets: np.ndarray, threads: int = 1, mem_gb: float = 16.0,
                devices=None, **options):
    """All reads against the adapter and its reverse complement on one or several GPUs (contiguous read ranges, one
    host thread + one handle per GPU, concatenated in read order; no collective)."""
    if devices is None:
        devices = [0]
    elif isinstance(devices, int):
        devices = list(range(devices))
    devices = list(devices)
    offsets = np.ascontiguousarray(offsets, dtype=np.int64)
    n = offsets.shape[0] - 1
    if len(devices) == 1:
        with Aligner(adapter, threads, mem_gb, devices[0], **options) as a:
            return a.align(read_bytes, offsets)
    bounds = shard_bounds(n, len(devices))
    results = [None] * len(devices)
    errors = []

    def work(i):
        try:
            lo, hi = bounds[i]
            with Aligner(adapter, threads, mem_gb / len(devices), devices[i], **options) as a:
                results[i] = a.align(read_bytes, offsets[lo:hi + 1])
        except Exception as e:
            errors.append(e)

    ts = [threading.Thread(target=work, args=(i,)) for i in range(len(devices))]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    if errors:
        raise errors[0]
    strand = np.concatenate([r[0] for r in results])
    dist = np.concatenate([r[1] for r in results])
    locs = np.concatenate([r[3] for r in results])
    offs, base = [], 0
    for r in results:
        offs.append(r[2][:-1] + base)
        base += int(r[2][-1])
    offs.append(np.array([base], dtype=np.int64))
    return strand, dist, np.concatenate(offs), locs


def measure_int32_peak(device: int = 0, mode: int = 0, repeats: int = 10) -> float:
    v = ctypes.c_double(0)
    _check(lib().sct_measure_int32_peak(device, mode, repeats, ctypes.byref(v)))
    return float(v.value)


def shard_bounds(n: int, parts: int):
    """Contiguous read ranges, one per GPU: [(first, last), ...]."""
    parts = max(1, parts)
    return [((n * i) // parts, (n * (i + 1)) // parts) for i in range(parts)]


def concat_results(parts):
    """Concatenate per-shard MatchResults in read order (the host-side 'gather'; no collective)."""
    from .match_trie import MatchResult
    if len(parts) == 1:
        return parts[0]
    min_ed = np.concatenate([p.min_ed for p in parts])
    n_ties = np.concatenate([p.n_ties for p in parts])
    ids = np.concatenate([p.tie_ids for p in parts])
    offs, base = [], 0
    for p in parts:
        offs.append(p.tie_offsets[:-1] + base)
        base += int(p.tie_offsets[-1])
    offs.append(np.array([base], dtype=np.int64))
    return MatchResult(min_ed, n_ties, np.concatenate(offs), ids)


class MatcherGroup:
    """One Matcher per GPU of this process (handles are created concurrently: CUDA context creation is the slow part and
    the CLI overlaps it with reading the -lr file).  Reads are split into contiguous ranges, one host thread per GPU
    (ctypes releases the GIL), results are concatenated in read order; no collective."""

    def __init__(self, barcodes: np.ndarray, max_error: int, threads: int = 16, mem_gb: float = 16.0, devices=None,
                 **options):
        if devices is None:
            devices = [0]
        elif isinstance(devices, int):
            devices = list(range(devices))
        self.devices = list(devices)
        g = len(self.devices)
        self.matchers = [None] * g
        errors = []

        def make(i):
            try:
                self.matchers[i] = Matcher(barcodes, max_error, threads, mem_gb / g, self.devices[i], **options)
            except Exception as e:      # re-raised in the caller
                errors.append(e)

        _run_threads(make, g)
        if errors:
            self.close()
            raise errors[0]

    def match(self, seg_bytes: np.ndarray, offsets: np.ndarray):
        offsets = np.ascontiguousarray(offsets, dtype=np.int64)
        n = offsets.shape[0] - 1
        g = len(self.matchers)
        if g == 1:
            return self.matchers[0].match(seg_bytes, offsets)
        bounds = shard_bounds(n, g)
        results = [None] * g
        errors = []

        def work(i):
            try:
                a, b = bounds[i]
                results[i] = self.matchers[i].match(seg_bytes, offsets[a:b + 1])
            except Exception as e:
                errors.append(e)

        _run_threads(work, g)
        if errors:
            raise errors[0]
        return concat_results(results)

    def close(self):
        for m in self.matchers:
            if m is not None:
                m.close()
        self.matchers = []

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def _run_threads(fn, n):
    if n == 1:
        fn(0)
        return
    ts = [threading.Thread(target=fn, args=(i,)) for i in range(n)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()


def match_segments(barcodes: np.ndarray, seg_bytes: np.ndarray, offsets: np.ndarray, max_error: int,
                   threads: int = 16, mem_gb: float = 16.0, devices=None, **options):
    """All segments against all barcodes on one or several GPUs of this process.
    devices: None = device 0; an int N = devices 0..N-1; or an explicit list."""
    with MatcherGroup(barcodes, max_error, threads, mem_gb, devices, **options) as g:
        return g.match(seg_bytes, offsets)


# ---- native host I/O (sct_io_*): parse / format / gzip with the host threads, no GPU involved ---------------------------
def host_threads(threads=None) -> int:
    n = os.cpu_count() or 1
    return max(1, min(int(threads) if threads else n, n, 64))


def _take_reads(h):
    """Copy the four arrays out of a sct_reads handle and free it -> (name_bytes, name_offsets, seg_bytes, seg_offsets)."""
    try:
        n = ctypes.c_int64(0)
        nb, sb = _u8p(), _u8p()
        no, so = _i64p(), _i64p()
        _check(lib().sct_io_reads_info(h, ctypes.byref(n), ctypes.byref(nb), ctypes.byref(no), ctypes.byref(sb),
                                       ctypes.byref(so)))
        n = int(n.value)
        name_off = np.ctypeslib.as_array(no, shape=(n + 1,)).copy()
        seg_off = np.ctypeslib.as_array(so, shape=(n + 1,)).copy()
        nbytes, sbytes = int(name_off[n]), int(seg_off[n])
        name_bytes = np.ctypeslib.as_array(nb, shape=(nbytes,)).copy() if nbytes else np.zeros(0, dtype=np.uint8)
        seg_bytes = np.ctypeslib.as_array(sb, shape=(sbytes,)).copy() if sbytes else np.zeros(0, dtype=np.uint8)
        return name_bytes, name_off, seg_bytes, seg_off
    finally:
        lib().sct_io_reads_free(h)


def io_read_long_reads(path: str, threads=None):
    """sct_io_read_long_reads -> (name_bytes u8, name_offsets i64[n+1], seg_bytes u8, seg_offsets i64[n+1])."""
    h = _vp()
    _check(lib().sct_io_read_long_reads(os.fsencode(path), host_threads(threads), ctypes.byref(h)))
    return _take_reads(h)


def io_read_fastqs(path: str, gzipped: bool = False, threads=None):
    """sct_io_read_fastqs, one file -> (name_bytes, name_offsets, read_bytes, read_offsets)."""
    h = _vp()
    _check(lib().sct_io_read_fastqs(os.fsencode(path), 1 if gzipped else 0, host_threads(threads), ctypes.byref(h)))
    return _take_reads(h)


def io_read_lr_bc_segments(path: str, threads=None):
    """sct_io_read_lr_bc_segments -> (seg_bytes, seg_offsets) of the rows whose location column is not 'NA'; a row that
    does not have exactly four columns raises ValueError like the reference's tuple unpacking (scTagger.py:820)."""
    h = _vp()
    try:
        _check(lib().sct_io_read_lr_bc_segments(os.fsencode(path), host_threads(threads), ctypes.byref(h)))
    except IndexError as e:
        raise ValueError(str(e)) from None
    _, _, sb, so = _take_reads(h)
    return sb, so


def io_read_whitelist(path: str, threads=None):
    """sct_io_read_whitelist -> (word_bytes, word_offsets): every line minus its last character (scTagger.py:798)."""
    h = _vp()
    _check(lib().sct_io_read_whitelist(os.fsencode(path), host_threads(threads), ctypes.byref(h)))
    _, _, wb, wo = _take_reads(h)
    return wb, wo


def _take_buffer(h):
    try:
        data = _u8p()
        size, rows = ctypes.c_int64(0), ctypes.c_int64(0)
        _check(lib().sct_io_buffer_info(h, ctypes.byref(data), ctypes.byref(size), ctypes.byref(rows)))
        return (ctypes.string_at(data, int(size.value)) if size.value else b""), int(rows.value)
    finally:
        lib().sct_io_buffer_free(h)


def _i64(a):
    a = np.ascontiguousarray(a, dtype=np.int64)
    return a if a.size else np.zeros(1, dtype=np.int64)


def _i32(a):
    a = np.ascontiguousarray(a, dtype=np.int32)
    return a if a.size else np.zeros(1, dtype=np.int32)


def io_format_matches(name_bytes, name_offsets, seg_bytes, seg_offsets, result, barcode_matrix, gzip_level=0,
                      threads=None):
    """sct_io_format_matches -> (bytes of the TSV rows, gzip members when gzip_level > 0; number of rows)."""
    bm = np.ascontiguousarray(barcode_matrix, dtype=np.uint8)
    n = int(np.asarray(seg_offsets).shape[0]) - 1
    nb, no, sb, so = _as_u8(name_bytes), _i64(name_offsets), _as_u8(seg_bytes), _i64(seg_offsets)
    me, nt, to, ti = _i32(result.min_ed), _i32(result.n_ties), _i64(result.tie_offsets), _i32(result.tie_ids)
    h = _vp()
    _check(lib().sct_io_format_matches(_ptr(nb, ctypes.c_uint8), _ptr(no, ctypes.c_int64), _ptr(sb, ctypes.c_uint8),
                                       _ptr(so, ctypes.c_int64), n, _ptr(me, ctypes.c_int32), _ptr(nt, ctypes.c_int32),
                                       _ptr(to, ctypes.c_int64), _ptr(ti, ctypes.c_int32),
                                       _ptr(_as_u8(bm), ctypes.c_uint8), int(bm.shape[0]), int(bm.shape[1]),
                                       int(gzip_level), host_threads(threads), ctypes.byref(h)))
    return _take_buffer(h)


def io_format_lr_bc(name_bytes, name_offsets, read_bytes, read_offsets, dist, loc, s, e, gzip_level=0, threads=None):
    """sct_io_format_lr_bc -> (bytes, rows)."""
    n = int(np.asarray(read_offsets).shape[0]) - 1
    nb, no, rb, ro = _as_u8(name_bytes), _i64(name_offsets), _as_u8(read_bytes), _i64(read_offsets)
    d, l, s_, e_ = _i64(dist), _i64(loc), _i64(s), _i64(e)
    h = _vp()
    _check(lib().sct_io_format_lr_bc(_ptr(nb, ctypes.c_uint8), _ptr(no, ctypes.c_int64), _ptr(rb, ctypes.c_uint8),
                                     _ptr(ro, ctypes.c_int64), n, _ptr(d, ctypes.c_int64), _ptr(l, ctypes.c_int64),
                                     _ptr(s_, ctypes.c_int64), _ptr(e_, ctypes.c_int64), int(gzip_level),
                                     host_threads(threads), ctypes.byref(h)))
    return _take_buffer(h)


def pack_names(names):
    """list of str/bytes -> (name_bytes u8, name_offsets i64[n+1])."""
    enc = [x.encode("latin-1") if isinstance(x, str) else bytes(x) for x in names]
    off = np.zeros(len(enc) + 1, dtype=np.int64)
    if enc:
        np.cumsum(np.fromiter((len(x) for x in enc), dtype=np.int64, count=len(enc)), out=off[1:])
    return np.frombuffer(b"".join(enc), dtype=np.uint8), off


WL_STAT_NAMES = ("launches", "chunks", "h2d_bytes", "d2h_bytes", "table_keys", "segments", "table_slots", "literal_words")


class Whitelist:
    """One handle = one GPU holding the whitelist hash table (sct_wl_create .. sct_wl_destroy): the Aho-Corasick
    automaton and Counter loop of extract_sr_bc_from_lr (reference scTagger.py:802-827)."""

    def __init__(self, word_bytes: np.ndarray, word_offsets: np.ndarray, threads: int = 16, device: int = 0):
        wb = _as_u8(word_bytes)
        wo = _i64(word_offsets)
        self.n_words = int(np.asarray(word_offsets).shape[0]) - 1
        self._h = _vp()
        _check(lib().sct_wl_create(int(device), _ptr(wb, ctypes.c_uint8), _ptr(wo, ctypes.c_int64), self.n_words,
                                   int(threads), ctypes.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib().sct_wl_destroy(self._h)
            self._h = _vp()

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def count(self, seg_bytes: np.ndarray, offsets: np.ndarray, per_segment: bool = False, out=None):
        """Adds one batch of segments to the counts; per_segment=True -> (n_hits int32[n], hit_word int32[n]), written
        into `out` = (n_hits, hit_word) when given.  Arrays in page-locked memory (`Pinned`) are copied from / to
        directly; pageable ones are staged by the library's host threads."""
        sb = _as_u8(seg_bytes)
        off = np.ascontiguousarray(offsets, dtype=np.int64)
        n = off.shape[0] - 1
        if per_segment:
            nh, hw = out if out is not None else (np.empty(max(n, 1), dtype=np.int32), np.empty(max(n, 1), dtype=np.int32))
            _check(lib().sct_wl_count(self._h, _ptr(sb, ctypes.c_uint8), _ptr(off, ctypes.c_int64), n,
                                      _ptr(nh, ctypes.c_int32), _ptr(hw, ctypes.c_int32)))
            return nh[:n], hw[:n]
        _check(lib().sct_wl_count(self._h, _ptr(sb, ctypes.c_uint8), _ptr(off, ctypes.c_int64), n, None, None))
        return None

    def result(self):
        """The Counter in insertion order -> (barcode index int32[u], count int64[u])."""
        n = ctypes.c_int64(0)
        idx, cnt = _i32p(), _i64p()
        _check(lib().sct_wl_result(self._h, ctypes.byref(n), ctypes.byref(idx), ctypes.byref(cnt)))
        u = int(n.value)
        if u == 0:
            return np.zeros(0, dtype=np.int32), np.zeros(0, dtype=np.int64)
        return np.ctypeslib.as_array(idx, shape=(u,)).copy(), np.ctypeslib.as_array(cnt, shape=(u,)).copy()

    def reset(self):
        _check(lib().sct_wl_reset(self._h))

    def stage(self, seg_bytes: np.ndarray, offsets: np.ndarray):
        sb = _as_u8(seg_bytes)
        off = np.ascontiguousarray(offsets, dtype=np.int64)
        _check(lib().sct_wl_stage(self._h, _ptr(sb, ctypes.c_uint8), _ptr(off, ctypes.c_int64), off.shape[0] - 1))

    def run_staged(self) -> float:
        ms = ctypes.c_float(0)
        _check(lib().sct_wl_run_staged(self._h, ctypes.byref(ms)))
        return float(ms.value)

    def stats(self) -> dict:
        s = np.zeros(8, dtype=np.int64)
        _check(lib().sct_wl_stats(self._h, _ptr(s, ctypes.c_int64)))
        return dict(zip(WL_STAT_NAMES, s.tolist()))
