"""Host-side mirror of the reference's `match_trie` stage (reference scTagger.py:602-640, 688-716, 758-790).

Same function names, argument meaning, file formats and error behaviour as the reference, so the
parity tests read like tests of the reference itself:

    read_sr_barcodes(path)        <- scTagger.py:631-640
    read_long_reads(path)         <- scTagger.py:602-628
    run_get_matches_prefix(...)   <- scTagger.py:688-716   (the seam: here it calls the CUDA library)
    match_trie(args)              <- scTagger.py:758-790

The compute is done by `libsctagger_b200.so` (C-ABI in include/sctagger_b200.h) through
`sctagger_b200._lib`; there is no CPU fallback — if the library or a GPU is missing the call raises.
"""
from __future__ import annotations

import gzip
import os
import sys
from dataclasses import dataclass

import numpy as np

_RC = np.arange(256, dtype=np.uint8)
for _a, _b in (("A", "T"), ("C", "G"), ("G", "C"), ("T", "A")):
    _RC[ord(_a)] = ord(_b)


def rev_compl(s: str) -> str:
    """Reverse complement with identity on non-ACGT characters (reference scTagger.py:148-156)."""
    b = np.frombuffer(s.encode("latin-1"), dtype=np.uint8)
    return _RC[b[::-1]].tobytes().decode("latin-1")


def prefix_k(threads: int) -> int:
    """Length k of the DNA prefixes the reference fans out over: smallest k with 4^k >= threads
    (reference scTagger.py:689-697).  k only matters for segments with non-ACGT bytes."""
    k, p = 0, 1
    while threads > p:
        p *= 4
        k += 1
    return k


def _read_all(path: str) -> bytes:
    if path.endswith(".gz"):
        with gzip.open(path, "rb") as f:
            return f.read()
    with open(path, "rb") as f:
        return f.read()


def _lines(data: bytes):
    lines = data.split(b"\n")
    if lines and lines[-1] == b"":
        lines.pop()
    return lines


def read_sr_barcodes(path: str):
    """Column 0 of every line, file order = barcode id (reference scTagger.py:631-640)."""
    print("Reading short reads barcodes")
    barcodes = [l.rstrip(b"\r").split(b"\t")[0].decode("latin-1") for l in _lines(_read_all(path))]
    print(f"There are {len(barcodes):,} SR barcodes")
    return barcodes


@dataclass
class LongReads:
    name_bytes: np.ndarray    # uint8, all read names concatenated (lr_names)
    name_offsets: np.ndarray  # int64[n+1]
    seg_bytes: np.ndarray     # uint8, all segments concatenated (lr_segs)
    offsets: np.ndarray       # int64[n+1] (lr_segs_idxs)

    def __len__(self):
        return int(self.offsets.shape[0]) - 1

    @property
    def names(self):
        """lr_names as a list of bytes (built on demand; the stage itself never needs it)."""
        raw, o = self.name_bytes.tobytes(), self.name_offsets.tolist()
        return [raw[o[i]:o[i + 1]] for i in range(len(self))]

    def segment(self, rid: int) -> bytes:
        return self.seg_bytes[self.offsets[rid]:self.offsets[rid + 1]].tobytes()


def read_long_reads(path: str, threads=None) -> LongReads:
    """Column 0 = read name, column 3 = segment (reference scTagger.py:602-628); gzip iff the name ends with '.gz'.
    Rows with fewer than four columns raise IndexError like the reference (scTagger.py:613-614).  Parsed natively
    (sct_io_read_long_reads: threaded, two memchr passes) instead of the reference's per-character Python loop."""
    from . import _lib
    print("Reading long reads barcodes")
    reads = LongReads(*_lib.io_read_long_reads(path, threads))
    print(f"There are {len(reads):,} LRs")
    return reads


@dataclass
class MatchResult:
    """Per read: minimum edit distance (-1 = no barcode within max_error -> no output row), number
    of tied patterns and the tied pattern ids `(bid << 1) | strand` (strand bit 1 = forward, 0 =
    reverse complement) ascending — i.e. the `sorted(bids)` order of reference scTagger.py:789."""
    min_ed: np.ndarray      # int32[n]
    n_ties: np.ndarray      # int32[n]
    tie_offsets: np.ndarray  # int64[n+1]
    tie_ids: np.ndarray     # int32[total]

    def ties(self, rid: int):
        return self.tie_ids[self.tie_offsets[rid]:self.tie_offsets[rid + 1]]

    def equals(self, other: "MatchResult") -> bool:
        return (np.array_equal(self.min_ed, other.min_ed) and np.array_equal(self.n_ties, other.n_ties)
                and np.array_equal(self.tie_offsets, other.tie_offsets) and np.array_equal(self.tie_ids, other.tie_ids))


def barcode_matrix(barcodes, barcode_length: int) -> np.ndarray:
    """(B, bl) uint8 matrix; asserts like the reference that every barcode has length -bl (scTagger.py:759-761)."""
    barcode_lens = {len(b) for b in barcodes}
    assert barcode_lens == {barcode_length}, barcode_lens
    raw = "".join(barcodes).encode("latin-1")
    return np.frombuffer(raw, dtype=np.uint8).reshape(len(barcodes), barcode_length).copy()


def run_get_matches_prefix(barcodes, reads: LongReads, max_error: int, threads: int, barcode_length: int,
                           mem_gb: float = 16.0, devices=None) -> MatchResult:
    """The seam the replacement cuts along (reference scTagger.py:688-716): all reads against all
    barcodes and their reverse complements, minimum edit distance and ties per read.
    `threads` keeps its reference meaning for results (it fixes k, see `prefix_k`) and is also the
    number of host worker threads; `mem_gb` (-m) sizes the streamed chunks; `devices` = GPUs to shard over."""
    from . import _lib
    bm = barcode_matrix(barcodes, barcode_length)
    return _lib.match_segments(bm, reads.seg_bytes, reads.offsets, max_error=max_error, threads=threads,
                               mem_gb=mem_gb, devices=devices)


def format_reads_rows(reads: LongReads, result: MatchResult, barcodes, gzip_level: int = 0, threads=None) -> bytes:
    """TSV rows exactly as the reference writes them (scTagger.py:777-789): only reads with a match, in read order;
    `name, e, n_ties, segment, comma-joined tied barcodes` (RC string for strand False).  gzip_level > 0 returns the
    same rows as gzip members.  Formatted natively (sct_io_format_matches)."""
    from . import _lib
    bl = len(barcodes[0]) if len(barcodes) else 1
    bm = barcode_matrix(barcodes, bl) if len(barcodes) else np.zeros((0, 1), dtype=np.uint8)
    data, _ = _lib.io_format_matches(reads.name_bytes, reads.name_offsets, reads.seg_bytes, reads.offsets, result, bm,
                                     gzip_level=gzip_level, threads=threads)
    return data


def format_rows(names, reads_seg_bytes: np.ndarray, offsets: np.ndarray, result: MatchResult, barcodes) -> bytes:
    """`format_reads_rows` for names given as a list (or a LongReads' `names`)."""
    from . import _lib
    nb, no = _lib.pack_names(names)
    return format_reads_rows(LongReads(nb, no, reads_seg_bytes, offsets), result, barcodes)


# `gzip.open` in the reference compresses at level 9 (30 s per 1 M rows on one core); level 4 costs 7 % more bytes and
# 0.25 s on 8 threads.  Only the decompressed bytes are comparable between writers.  SCT_GZIP_LEVEL=1..9 overrides.
GZIP_LEVEL = min(9, max(1, int(os.environ.get("SCT_GZIP_LEVEL", "4"))))


def write_matches(outfile, data: bytes):
    """Plain bytes `data` -> gzip iff the name ends with 'gz', plain file otherwise, STDOUT when absent
    (scTagger.py:770-776)."""
    if outfile:
        if outfile.endswith("gz"):
            with gzip.open(outfile, "wb") as f:
                f.write(data)
        else:
            with open(outfile, "wb") as f:
                f.write(data)
    else:
        sys.stdout.buffer.write(data)
        sys.stdout.buffer.flush()


def _match_trie_result(args):
    import threading
    from . import _lib
    barcodes = read_sr_barcodes(args.short_read_barcodes)
    barcode_lens = {len(b) for b in barcodes}
    assert barcode_lens == {args.barcode_length}, barcode_lens
    # the GPU handles (CUDA context + barcode index) are created while the host parses the -lr file
    box = {}

    def make_group():
        try:
            box["group"] = _lib.MatcherGroup(barcode_matrix(barcodes, args.barcode_length), args.max_error, args.threads,
                                             args.mem, getattr(args, "devices", None), reserve_segments=1 << 20)
        except Exception as e:
            box["error"] = e

    t = threading.Thread(target=make_group)
    t.start()
    try:
        reads = read_long_reads(args.long_read_segments, args.threads)
    except BaseException:
        t.join()
        if "group" in box:
            box["group"].close()
        raise
    t.join()
    if "error" in box:
        raise box["error"]
    with box["group"] as group:
        result = group.match(reads.seg_bytes, reads.offsets)
    return barcodes, reads, result


def match_trie_bytes(args) -> bytes:
    """Everything of reference scTagger.py:758-790 except opening the output file: -> the TSV rows."""
    barcodes, reads, result = _match_trie_result(args)
    return format_reads_rows(reads, result, barcodes)


def match_trie(args):
    """Stage driver, same steps as reference scTagger.py:758-790.  With a `gz` output name the rows are compressed by
    the native writer (parallel gzip members) and written as they are."""
    barcodes, reads, result = _match_trie_result(args)
    if args.outfile and args.outfile.endswith("gz"):
        data = format_reads_rows(reads, result, barcodes, gzip_level=GZIP_LEVEL)
        with open(args.outfile, "wb") as f:
            f.write(data)
    else:
        write_matches(args.outfile, format_reads_rows(reads, result, barcodes))
