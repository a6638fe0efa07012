"""TEST INFRASTRUCTURE ONLY — ctypes binding of the CPU oracle (oracle/*.c).

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs (`cpu_baseline`, `--impl reference`) may import
this module; nothing under sctagger_b200/ does.  `build()` compiles oracle/_build/libsct_oracle.so with the
Makefile next to this file (gcc + OpenMP); the prebuilt .so travels to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libsct_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith(".c")]
    stale = (not os.path.exists(_SO)) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs)
    if force or stale:
        subprocess.run(["make", "-C", _HERE] + (["-B"] if force else []), check=True,
                       stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
        u8p = ctypes.POINTER(ctypes.c_uint8)
        i32p = ctypes.POINTER(ctypes.c_int32)
        i64p = ctypes.POINTER(ctypes.c_int64)
        for name in ("sct_oracle_trie", "sct_oracle_anchored", "sct_oracle_dp"):
            fn = getattr(_lib, name)
            fn.restype = ctypes.c_int
            fn.argtypes = [u8p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                           u8p, i64p, ctypes.c_int64, i32p, i32p, i64p, ctypes.POINTER(i32p)]
        _lib.sct_oracle_free.argtypes = [ctypes.c_void_p]
        _lib.sct_oracle_free.restype = None
    return _lib


def _ptr(a, ct):
    return a.ctypes.data_as(ctypes.POINTER(ct))


def match(barcodes: np.ndarray, seg_bytes: np.ndarray, offsets: np.ndarray, max_error: int, threads: int = 16,
          workers: int = 0, algo: str = "trie"):
    """Run one of the oracle restatements.  barcodes: (B, bl) uint8; seg_bytes uint8; offsets int64[n+1].
    Returns (min_ed int32[n] (-1 = no row), n_ties int32[n], tie_offsets int64[n+1], tie_ids int32[total]) with
    tie ids `(bid << 1) | strand` ascending, strand bit 1 = forward (reference scTagger.py:789 order)."""
    L = lib()
    fn = {"trie": L.sct_oracle_trie, "anchored": L.sct_oracle_anchored, "dp": L.sct_oracle_dp}[algo]
    barcodes = np.ascontiguousarray(barcodes, dtype=np.uint8)
    B, bl = barcodes.shape
    seg_bytes = np.ascontiguousarray(seg_bytes, dtype=np.uint8)
    if seg_bytes.size == 0:
        seg_bytes = np.zeros(1, dtype=np.uint8)
    offsets = np.ascontiguousarray(offsets, dtype=np.int64)
    n = offsets.shape[0] - 1
    min_ed = np.empty(max(n, 1), dtype=np.int32)
    n_ties = np.empty(max(n, 1), dtype=np.int32)
    tie_off = np.zeros(n + 1, dtype=np.int64)
    ids = ctypes.POINTER(ctypes.c_int32)()
    rc = fn(_ptr(barcodes, ctypes.c_uint8), B, bl, max_error, threads, workers, _ptr(seg_bytes, ctypes.c_uint8),
            _ptr(offsets, ctypes.c_int64), n, _ptr(min_ed, ctypes.c_int32), _ptr(n_ties, ctypes.c_int32),
            _ptr(tie_off, ctypes.c_int64), ctypes.byref(ids))
    if rc != 0:
        raise RuntimeError(f"oracle {algo} failed with code {rc}")
    total = int(tie_off[n])
    tie_ids = np.ctypeslib.as_array(ids, shape=(max(total, 1),))[:total].copy()
    L.sct_oracle_free(ids)
    return min_ed[:n], n_ties[:n], tie_off, tie_ids


def align(adapter: bytes, read_bytes: np.ndarray, offsets: np.ndarray, workers: int = 0, algo: str = "dp"):
    """Adapter alignment oracle (oracle/align_oracle.c; PARITY UNPINNED, see its header): what the reference's
    get_alns (scTagger.py:176-196) returns for every read, as arrays:
    (strand int8[n] ('+', '-' as ASCII, 0 = 'NA'), dist int32[n] (-1 = NA), loc_offsets int64[n+1], locs int32[total])."""
    L = lib()
    fn = {"dp": L.sct_oracle_align, "bv": L.sct_oracle_align_bv}[algo]
    u8p = ctypes.POINTER(ctypes.c_uint8)
    i32p = ctypes.POINTER(ctypes.c_int32)
    fn.restype = ctypes.c_int
    fn.argtypes = [u8p, ctypes.c_int, u8p, ctypes.POINTER(ctypes.c_int64), ctypes.c_int64, ctypes.c_int,
                   ctypes.POINTER(ctypes.c_int8), i32p, ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(i32p)]
    ad = np.frombuffer(bytes(adapter), dtype=np.uint8).copy()
    read_bytes = np.ascontiguousarray(read_bytes, dtype=np.uint8)
    if read_bytes.size == 0:
        read_bytes = np.zeros(1, dtype=np.uint8)
    offsets = np.ascontiguousarray(offsets, dtype=np.int64)
    n = offsets.shape[0] - 1
    strand = np.zeros(max(n, 1), dtype=np.int8)
    dist = np.empty(max(n, 1), dtype=np.int32)
    loc_off = np.zeros(n + 1, dtype=np.int64)
    ids = i32p()
    rc = fn(_ptr(ad, ctypes.c_uint8), int(ad.size), _ptr(read_bytes, ctypes.c_uint8), _ptr(offsets, ctypes.c_int64), n,
            int(workers), _ptr(strand, ctypes.c_int8), _ptr(dist, ctypes.c_int32), _ptr(loc_off, ctypes.c_int64),
            ctypes.byref(ids))
    if rc != 0:
        raise RuntimeError(f"align oracle failed with code {rc}")
    total = int(loc_off[n])
    locs = np.ctypeslib.as_array(ids, shape=(max(total, 1),))[:total].copy()
    L.sct_oracle_free(ids)
    return strand[:n], dist[:n], loc_off, locs


def hw_locations(query: bytes, target: bytes, cap: int = 256):
    """edlib.align(query, target, mode='HW', task='locations') as restated by oracle/align_oracle.c:
    -> (editDistance, [(start, end), ...]).  Used to check the restatement against the vectors edlib publishes."""
    L = lib()
    fn = L.sct_oracle_hw_locations
    u8p = ctypes.POINTER(ctypes.c_uint8)
    i32p = ctypes.POINTER(ctypes.c_int32)
    fn.restype = ctypes.c_int
    fn.argtypes = [u8p, ctypes.c_int, u8p, ctypes.c_int64, i32p, i32p, i32p, ctypes.c_int]
    q = np.frombuffer(bytes(query), dtype=np.uint8).copy()
    t = np.frombuffer(bytes(target) or b"\0", dtype=np.uint8).copy()
    d = ctypes.c_int32(0)
    starts, ends = np.zeros(cap, np.int32), np.zeros(cap, np.int32)
    k = fn(_ptr(q, ctypes.c_uint8), int(q.size), _ptr(t, ctypes.c_uint8), len(target), ctypes.byref(d),
           _ptr(starts, ctypes.c_int32), _ptr(ends, ctypes.c_int32), cap)
    return int(d.value), [(int(starts[i]), int(ends[i])) for i in range(k)]
