"""Host-side mirror of the reference's `extract_lr_bc` stage (reference scTagger.py:159-320, 362-383).

Same function names, argument meaning, file formats and (stderr) messages as the reference:

    read_fastqs(fastqs, gzipped)          <- scTagger.py:159-173
    get_lr_bc_alns(seqs, threads)         <- scTagger.py:307-314   (the seam: here it calls the CUDA library)
    get_ranges / get_possible_ranges      <- scTagger.py:199-252
    set_global_range_dicts                <- scTagger.py:255-264
    filter_aln / get_filter_alns          <- scTagger.py:277-304
    output_matches                        <- scTagger.py:317-320
    extract_lr_bc(args)                   <- scTagger.py:362-383

The two edlib alignments per read (scTagger.py:180-181) run on the GPU through `sct_align` (include/sctagger_b200.h);
there is no CPU fallback.  Reads are kept as one concatenated byte array + offsets instead of a list of Python strings,
alignments as arrays (`Alignments`) instead of a list of tuples; `Alignments.tuples()` gives the reference's view.
"""
from __future__ import annotations

import sys
from collections import deque
from dataclasses import dataclass

import numpy as np

from .match_trie import GZIP_LEVEL, rev_compl

DEFAULT_ADAPTER = "CTACACGACGCTCTTCCGATCT"      # scTagger.py:42


@dataclass
class Reads:
    name_bytes: np.ndarray    # uint8, all rnames concatenated
    name_offsets: np.ndarray  # int64[n+1]
    read_bytes: np.ndarray    # uint8, all sequences concatenated
    offsets: np.ndarray       # int64[n+1]

    def __len__(self):
        return int(self.offsets.shape[0]) - 1

    @property
    def names(self):
        raw, o = self.name_bytes.tobytes(), self.name_offsets.tolist()
        return [raw[o[i]:o[i + 1]] for i in range(len(self))]

    def seq(self, i: int) -> bytes:
        return self.read_bytes[self.offsets[i]:self.offsets[i + 1]].tobytes()


@dataclass
class Alignments:
    """What get_alns returns for every read (scTagger.py:192-196), as arrays."""
    strand: np.ndarray       # int8: ord('+'), ord('-') or 0 for 'NA'
    dist: np.ndarray         # int32, -1 for 'NA'
    loc_offsets: np.ndarray  # int64[n+1]
    locs: np.ndarray         # int32

    def __len__(self):
        return self.dist.shape[0]

    def tuples(self):
        """The reference's list of (s, d, locs) tuples ('NA', -1, ['NA'] when the two distances tie)."""
        out = []
        lo = self.loc_offsets
        for i in range(len(self)):
            if self.strand[i] == 0:
                out.append(("NA", -1, ["NA"]))
            else:
                out.append((chr(self.strand[i]), int(self.dist[i]), self.locs[lo[i]:lo[i + 1]].tolist()))
        return out


def read_fastqs(fastqs, gzipped=False, threads=None) -> Reads:
    """Line 0 of every record: name = first whitespace-separated token minus its first character; line 1: the sequence
    with trailing whitespace stripped (scTagger.py:159-173).  gzip iff `gzipped` or the name ends with '.gz'.  Parsed
    natively (sct_io_read_fastqs), one file at a time, concatenated in argument order."""
    from . import _lib
    parts = []
    for fastq in fastqs:
        print(f"Reading {fastq}", file=sys.stderr)
        parts.append(_lib.io_read_fastqs(fastq, bool(gzipped), threads))
    if len(parts) == 1:
        return Reads(*parts[0])
    if not parts:
        z8, z64 = np.zeros(0, dtype=np.uint8), np.zeros(1, dtype=np.int64)
        return Reads(z8, z64, z8, z64.copy())

    def cat_off(k):
        out, base = [], 0
        for p in parts:
            out.append(p[k][:-1] + base)
            base += int(p[k][-1])
        out.append(np.array([base], dtype=np.int64))
        return np.concatenate(out)

    return Reads(np.concatenate([p[0] for p in parts]), cat_off(1), np.concatenate([p[2] for p in parts]), cat_off(3))


def get_lr_bc_alns(reads: Reads, threads: int, adapter: str = DEFAULT_ADAPTER, mem_gb: float = 16.0,
                   devices=None) -> Alignments:
    """The seam (scTagger.py:307-314 mapping get_alns 176-196 over the reads): adapter `a1` and `a2 = rev_compl(a1)`
    against every whole read, semi-global; strand, distance and locations per read."""
    from . import _lib
    print(f"Aligning {adapter} to {len(reads)} reads on {threads} threads", file=sys.stderr)
    s, d, lo, locs = _lib.align_reads(adapter.encode("latin-1"), reads.read_bytes, reads.offsets, threads=threads,
                                      mem_gb=mem_gb, devices=devices)
    return Alignments(s, d, lo, locs)


def get_ranges(data):
    """Peak picking on the histogram of adapter locations (scTagger.py:199-233): repeatedly take the highest bin, stop
    when fewer than 1 % of all locations lie within 20 bins of it, otherwise grow the peak left and right over bins
    holding more than 0.1 % of the locations, report (first, last) and clear 20 bins of margin around it."""
    ranges = []
    data = np.asarray(data, dtype=np.int64)
    if data.size == 0:
        return ranges
    min_l, max_l = int(data.min()), int(data.max())
    F = np.bincount(data - min_l, minlength=max_l - min_l + 1).astype(np.float64)
    T = F.sum()
    n = F.shape[0]
    while True:
        peak = int(np.argmax(F))
        around = F[max(0, peak - 20):min(peak + 20, n)].sum()
        print(f"--> {around / T: 5.2%} of strend reads fall around {min_l + peak}", file=sys.stderr)
        if around < 0.01 * T:
            break
        first = last = peak
        queue = deque([peak])
        while queue:
            i = queue.popleft()
            F[i] = 0
            if i <= peak and i - 1 > 0 and F[i - 1] > T * 0.001:      # note: bin 0 is never joined (i - 1 > 0)
                queue.append(i - 1)
                first = i - 1
            if i >= peak and i + 1 < n and F[i + 1] > T * 0.001:
                queue.append(i + 1)
                last = i + 1
        F[max(0, first - 20):min(last + 20, n)] = 0
        ranges.append((min_l + first, min_l + last))
    return ranges


def get_possible_ranges(alns: Alignments):
    """Locations of all reads aligned with 0 <= d <= 5, per strand (scTagger.py:236-252)."""
    ok = (alns.dist >= 0) & (alns.dist <= 5)
    per_loc_read = np.repeat(np.arange(len(alns)), np.diff(alns.loc_offsets))
    keep = ok[per_loc_read]
    plus = alns.strand[per_loc_read] == ord("+")
    ranges_f = get_ranges(alns.locs[keep & plus])
    print(f"Found these ranges on + strand:\t{ranges_f}", file=sys.stderr)
    ranges_r = get_ranges(alns.locs[keep & ~plus])
    print(f"Found these ranges on - strand:\t{ranges_r}", file=sys.stderr)
    return ranges_f, ranges_r


def parse_ranges_flag(ranges):
    """-g f20:40 r1:30 ... (scTagger.py:106-126): 1-based inclusive intervals -> per strand [(s, e)] whose dictionary
    covers s..e INCLUSIVE ('f': s-1..e;  'r': -e..-s+1)."""
    out = [[], []]
    for r in ranges:
        assert r[0] in "fr", r
        strand = r[0]
        r = r[1:].split(":")
        assert len(r) == 2, r
        s, e = int(r[0]), int(r[1])
        assert 0 < s <= e, (s, e)
        if strand == "f":
            idx, s, e = 0, s - 1, e
        else:
            idx, s, e = 1, -e, -s + 1
        out[idx].append((s, e))
    return out


class RangeDicts:
    """The reference's `ranges_dicts` (location -> index of the range holding it), as half-open intervals.
    Ranges from -g cover s..e inclusive (scTagger.py:120); detected ranges cover s..e-1 (scTagger.py:261) — kept as is."""

    def __init__(self, ranges, inclusive_end: bool):
        self.intervals = []
        for R in ranges:
            iv = [(int(s), int(e) + (1 if inclusive_end else 0)) for s, e in R]
            seen = {}
            for idx, (s, e) in enumerate(iv):
                for i in range(s, e):
                    assert i not in seen, (i, seen)
                    seen[i] = idx
            self.intervals.append(iv)

    def lookup(self, range_idx: int, locs: np.ndarray) -> np.ndarray:
        hit = np.full(locs.shape, -1, dtype=np.int64)
        for idx, (s, e) in enumerate(self.intervals[range_idx]):
            hit[(locs >= s) & (locs < e)] = idx
        return hit


def set_global_range_dicts(ranges) -> RangeDicts:
    return RangeDicts(ranges, inclusive_end=False)


def get_filter_alns(alns: Alignments, range_dicts: RangeDicts, num_bp_after: int, threads: int = 1):
    """filter_aln for every read (scTagger.py:277-304): an alignment survives iff all its locations fall into exactly
    one range of its strand; -> (dist int64[n] (-1 = filtered), loc int64[n] (valid where dist >= 0), s, e)."""
    print(f"Filtering alignments using ranges on {threads} threads", file=sys.stderr)
    n = len(alns)
    lo = alns.loc_offsets
    cnt = np.diff(lo)
    per_loc_read = np.repeat(np.arange(n), cnt)
    minus_read = alns.strand == ord("-")
    hit = np.where(minus_read[per_loc_read], range_dicts.lookup(1, alns.locs), range_dicts.lookup(0, alns.locs))
    has = cnt > 0
    starts = lo[:-1][has]
    hmin = np.full(n, -1, dtype=np.int64)
    hmax = np.full(n, -1, dtype=np.int64)
    lmin = np.zeros(n, dtype=np.int64)
    lmax = np.zeros(n, dtype=np.int64)
    if starts.size:
        hmin[has] = np.minimum.reduceat(hit, starts)
        hmax[has] = np.maximum.reduceat(hit, starts)
        lmin[has] = np.minimum.reduceat(alns.locs.astype(np.int64), starts)
        lmax[has] = np.maximum.reduceat(alns.locs.astype(np.int64), starts)
    ok = has & (hmin >= 0) & (hmin == hmax) & (alns.strand != 0)
    dist = np.where(ok, alns.dist.astype(np.int64), -1)
    s = np.full(n, -1, dtype=np.int64)
    e = np.full(n, -1, dtype=np.int64)
    plus = ok & ~minus_read
    minus = ok & minus_read
    s[plus] = np.maximum(0, lmin[plus] - 2)
    e[plus] = lmax[plus] + num_bp_after
    s[minus] = lmin[minus] - num_bp_after
    e[minus] = np.minimum(0, lmax[minus] + 2)
    loc = np.where(plus, s, e)
    return dist, loc, s, e


def format_rows(reads: Reads, filtered, gzip_level: int = 0, threads=None) -> bytes:
    """`rname \\t dist \\t loc \\t seq[s:e or None]` per read (scTagger.py:317-320), loc = 'NA' for filtered reads.
    Formatted natively (sct_io_format_lr_bc); gzip_level > 0 returns the rows as gzip members."""
    from . import _lib
    dist, loc, s, e = filtered
    data, _ = _lib.io_format_lr_bc(reads.name_bytes, reads.name_offsets, reads.read_bytes, reads.offsets, dist, loc, s,
                                   e, gzip_level=gzip_level, threads=threads)
    return data


def output_matches(reads: Reads, filtered, outfile, threads=None):
    """gzip when -o is given (whatever its suffix, scTagger.py:377-378), STDOUT otherwise."""
    print(f"Writng to {outfile}", file=sys.stderr)
    if outfile:
        with open(outfile, "wb") as f:
            f.write(format_rows(reads, filtered, gzip_level=GZIP_LEVEL, threads=threads))
    else:
        sys.stdout.buffer.write(format_rows(reads, filtered, threads=threads))
        sys.stdout.buffer.flush()


def extract_lr_bc(args):
    """Stage driver, same steps as reference scTagger.py:362-383 (plotting, -p, is not part of this path)."""
    reads = read_fastqs(args.reads, args.gzipped, args.threads)
    alns = get_lr_bc_alns(reads, args.threads, adapter=args.short_read_adapter,
                          devices=getattr(args, "devices", None))
    if len(args.ranges[0]) + len(args.ranges[1]) == 0:
        print("No ranges for SR adapters have been preset. Detecting directly from data...", file=sys.stderr)
        args.ranges = get_possible_ranges(alns)
        range_dicts = set_global_range_dicts(args.ranges)
    else:
        range_dicts = RangeDicts(args.ranges, inclusive_end=True)
    filtered = get_filter_alns(alns, range_dicts, args.num_bp_after, args.threads)
    output_matches(reads, filtered, args.outfile, args.threads)
    return alns, filtered
