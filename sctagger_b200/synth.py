"""Seeded synthetic inputs for the `match_trie` hot path.

Barcodes: B distinct 16-mers, i.i.d. uniform over ACGT, written `barcode\\tcount`.
Segments: rows `read{i}\\t{d}\\t{loc}\\t{seg}` in the format `extract_lr_bc` writes
(reference scTagger.py:317-320) and `read_long_reads` parses (scTagger.py:602-628).

Everything is vectorised numpy so that 10 M segments are generated in seconds; the
generator is deterministic in (seed, n, B).  Two output forms:
  * text files (gz or plain) for the CLI / reference harness,
  * in-memory arrays (concatenated segment bytes + int32/int64 offsets) for the C-ABI.
"""
from __future__ import annotations

import gzip
import numpy as np

ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
_COMP = np.arange(256, dtype=np.uint8)
_COMP[ord("A")], _COMP[ord("C")], _COMP[ord("G")], _COMP[ord("T")] = ord("T"), ord("G"), ord("C"), ord("A")

# seeds fixed by SURVEY.md §8(d)
SEEDS = {"cfg1": 1001, "cfg2": 1002, "cfg3": 1003, "cfg4": 1003, "cfg5": 1005}


def make_barcodes(n_barcodes: int, seed: int, bl: int = 16) -> np.ndarray:
    """(B, bl) uint8 ASCII matrix of distinct random barcodes, generation order = file order = bid."""
    rng = np.random.default_rng(seed)
    out = np.empty((0, bl), dtype=np.uint8)
    seen = set()
    while out.shape[0] < n_barcodes:
        need = n_barcodes - out.shape[0]
        cand = ACGT[rng.integers(0, 4, size=(need + 16, bl))]
        keep = []
        for row in cand:
            k = row.tobytes()
            if k not in seen:
                seen.add(k)
                keep.append(row)
                if len(keep) == need:
                    break
        if keep:
            out = np.concatenate([out, np.stack(keep)], axis=0)
    return out


def barcode_counts(n_barcodes: int, seed: int) -> np.ndarray:
    rng = np.random.default_rng(seed + 7919)
    return rng.integers(1, 1001, size=n_barcodes)


def make_segments(barcodes: np.ndarray, n_segments: int, seed: int):
    """Return (seg_bytes uint8[sum len], offsets int64[n+1]).

    seg = L(3 random) + core + R(random) ; core = mutated barcode (p=.9) or random 16-mer (p=.1);
    total length 22 (p=.9) or U[23,28] (p=.1); reverse-complemented with p=.5.  ACGT only.
    """
    rng = np.random.default_rng(seed + 104729)
    B, bl = barcodes.shape
    n = n_segments
    from_bc = rng.random(n) < 0.9
    which = rng.integers(0, B, size=n)
    n_edits = rng.choice(5, size=n, p=[0.50, 0.25, 0.12, 0.08, 0.05])
    n_edits[~from_bc] = 0
    tot_len = np.where(rng.random(n) < 0.9, 22, rng.integers(23, 29, size=n)).astype(np.int64)
    do_rc = rng.random(n) < 0.5

    # core as a padded matrix (bl + 4 wide: up to 4 insertions)
    W = bl + 4
    core = np.zeros((n, W), dtype=np.uint8)
    core[:, :bl] = barcodes[which]
    rnd_core = ACGT[rng.integers(0, 4, size=(n, bl))]
    core[~from_bc, :bl] = rnd_core[~from_bc]
    core_len = np.full(n, bl, dtype=np.int64)
    cols = np.arange(W)[None, :]
    for e in range(4):
        act = n_edits > e
        kind = rng.integers(0, 3, size=n)  # 0 sub, 1 ins, 2 del
        u = rng.random(n)
        newb = ACGT[rng.integers(0, 4, size=n)]
        # substitution at pos in [0, len)
        m = act & (kind == 0)
        pos = np.minimum((u * core_len).astype(np.int64), core_len - 1)
        idx = np.nonzero(m)[0]
        core[idx, pos[idx]] = newb[idx]
        # insertion before pos in [0, len]
        m = act & (kind == 1)
        idx = np.nonzero(m)[0]
        if idx.size:
            p = np.minimum((u[idx] * (core_len[idx] + 1)).astype(np.int64), core_len[idx])
            sub = core[idx]
            src = np.where(cols > p[:, None], cols - 1, cols)
            sub = np.take_along_axis(sub, src, axis=1)
            sub[np.arange(idx.size), p] = newb[idx]
            core[idx] = sub
            core_len[idx] += 1
        # deletion at pos in [0, len)
        m = act & (kind == 2) & (core_len > 1)
        idx = np.nonzero(m)[0]
        if idx.size:
            p = np.minimum((u[idx] * core_len[idx]).astype(np.int64), core_len[idx] - 1)
            sub = core[idx]
            src = np.where(cols >= p[:, None], np.minimum(cols + 1, W - 1), cols)
            sub = np.take_along_axis(sub, src, axis=1)
            core[idx] = sub
            core_len[idx] -= 1

    tot_len = np.maximum(tot_len, core_len + 3)
    Lmax = int(tot_len.max())
    seg = ACGT[rng.integers(0, 4, size=(n, Lmax))]  # random fill: L and R come from here
    # place the core at offset 3
    c = np.arange(Lmax)[None, :]
    in_core = (c >= 3) & (c < 3 + core_len[:, None])
    core_pad = np.zeros((n, Lmax), dtype=np.uint8)
    core_pad[:, 3:3 + W] = core[:, : max(0, min(W, Lmax - 3))]
    seg = np.where(in_core, core_pad, seg)
    # reverse-complement in place (within each row's own length)
    ridx = np.nonzero(do_rc)[0]
    if ridx.size:
        sub = seg[ridx]
        ln = tot_len[ridx]
        src = np.clip(ln[:, None] - 1 - c, 0, Lmax - 1)
        sub = _COMP[np.take_along_axis(sub, src, axis=1)]
        seg[ridx] = sub
    valid = c < tot_len[:, None]
    seg_bytes = seg[valid]  # row-major flatten keeps per-row order
    offsets = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(tot_len, out=offsets[1:])
    return np.ascontiguousarray(seg_bytes), offsets


def _open_w(path):
    return gzip.open(path, "wb", compresslevel=1) if str(path).endswith(".gz") else open(path, "wb")


def write_barcodes(path, barcodes: np.ndarray, seed: int):
    counts = barcode_counts(barcodes.shape[0], seed)
    with _open_w(path) as f:
        f.write(b"".join(row.tobytes() + b"\t%d\n" % c for row, c in zip(barcodes, counts)))


def read_names(n: int, first_index: int = 0):
    """`read{i}` for i = first_index .. first_index+n-1 as (name_bytes, name_offsets)."""
    a = np.char.add(b"read", np.arange(first_index, first_index + n).astype("S20"))
    u8 = a.view(np.uint8).reshape(n, -1) if n else np.zeros((0, 1), dtype=np.uint8)
    mask = u8 != 0
    off = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(mask.sum(axis=1), out=off[1:])
    return np.ascontiguousarray(u8[mask]), off


def write_segments(path, seg_bytes: np.ndarray, offsets: np.ndarray, first_index: int = 0):
    """`read{i} \\t 0 \\t 0 \\t segment` rows — the -lr format `output_matches` of extract_lr_bc writes
    (scTagger.py:317-320) — through the native writer (gzip members when the name ends with .gz)."""
    from . import _lib
    n = offsets.shape[0] - 1
    nb, no = read_names(n, first_index)
    z = np.zeros(max(n, 1), dtype=np.int64)
    data, _ = _lib.io_format_lr_bc(nb, no, seg_bytes, offsets, z, z, z, z,
                                   gzip_level=1 if str(path).endswith(".gz") else 0)
    with open(path, "wb") as f:
        f.write(data)


ADAPTER = b"CTACACGACGCTCTTCCGATCT"   # reference default -sa (scTagger.py:42)


def make_reads(n_reads: int, seed: int, adapter: bytes = ADAPTER, median_len: int = 900, min_len: int = 200,
               max_len: int = 5000):
    """BASELINE config 5: long reads with log-normal lengths (median 900, clipped 200..5000), uniform
    ACGT; 85 % carry the adapter with 8 % per-base errors — forward near the 5' end (offset U[10,60]) or reverse-
    complemented near the 3' end, 50/50 — followed by 16 + 12 random bases and T x 30.
    Returns (read_bytes uint8[sum len], offsets int64[n+1])."""
    rng = np.random.default_rng(seed + 15485863)
    lens = np.clip(np.exp(rng.normal(np.log(median_len), 0.6, size=n_reads)), min_len, max_len).astype(np.int64)
    offsets = np.zeros(n_reads + 1, dtype=np.int64)
    np.cumsum(lens, out=offsets[1:])
    out = ACGT[rng.integers(0, 4, size=int(offsets[-1]))]
    ad = np.frombuffer(adapter, dtype=np.uint8)
    m = ad.size
    has = rng.random(n_reads) < 0.85
    rev = rng.random(n_reads) < 0.5
    off5 = rng.integers(10, 61, size=n_reads)
    W = m + 8                                   # room for insertions
    for r in np.nonzero(has)[0]:
        # mutate the adapter: per-base sub / ins / del with total rate 8 %
        a = []
        for ch in ad:
            u = rng.random()
            if u < 0.08 / 3:
                a.append(int(ACGT[rng.integers(0, 4)]))
            elif u < 2 * 0.08 / 3:
                a.append(int(ch))
                a.append(int(ACGT[rng.integers(0, 4)]))
            elif u < 0.08:
                continue
            else:
                a.append(int(ch))
        a = a[:W]
        block = np.concatenate([np.array(a, dtype=np.uint8), ACGT[rng.integers(0, 4, size=28)],
                                np.full(30, ord("T"), dtype=np.uint8)])
        L = int(lens[r])
        block = block[: max(0, L - int(off5[r]))]
        if rev[r]:
            block = _COMP[block[::-1]]
            s = int(offsets[r + 1]) - int(off5[r]) - block.size
        else:
            s = int(offsets[r]) + int(off5[r])
        out[s:s + block.size] = block
    return out, offsets


def write_fastq(path, read_bytes: np.ndarray, offsets: np.ndarray, first_index: int = 0):
    raw = read_bytes.tobytes()
    with _open_w(path) as f:
        buf = []
        for i in range(offsets.shape[0] - 1):
            s = raw[offsets[i]:offsets[i + 1]]
            buf.append(b"@read%d some description\n%s\n+\n%s\n" % (first_index + i, s, b"I" * len(s)))
            if len(buf) >= 16384:
                f.write(b"".join(buf))
                buf = []
        f.write(b"".join(buf))


# ---- BASELINE configs[2]/[3]: ONE 10 M-segment input, defined block by block so that a rank can build just its shard ----
BLOCK_SEGMENTS = 250_000
BLOCK_SEED0 = 5000


def concat_segments(parts):
    """[(seg_bytes, offsets), ...] -> one (seg_bytes, offsets)."""
    if len(parts) == 1:
        return parts[0]
    sb = np.concatenate([p[0] for p in parts])
    offs, base = [np.zeros(1, dtype=np.int64)], 0
    for p in parts:
        offs.append(p[1][1:] + base)
        base += int(p[1][-1])
    return sb, np.concatenate(offs)


def make_segments_range(barcodes: np.ndarray, first: int, last: int, block: int = BLOCK_SEGMENTS, seed0: int = BLOCK_SEED0):
    """Segments [first, last) of the blocked workload: block b = make_segments(barcodes, block, seed0 + b).  The input is
    the same for every sharding (rank r of N builds only the blocks its read range touches)."""
    parts = []
    for b in range(first // block, (max(last, first + 1) - 1) // block + 1):
        sb, off = make_segments(barcodes, block, seed0 + b)
        lo, hi = max(first, b * block) - b * block, min(last, (b + 1) * block) - b * block
        if hi > lo:
            parts.append((sb[off[lo]:off[hi]], off[lo:hi + 1] - off[lo]))
    if not parts:
        return np.zeros(0, dtype=np.uint8), np.zeros(1, dtype=np.int64)
    return concat_segments(parts)


def lengthen(seg_bytes: np.ndarray, offsets: np.ndarray, extra: int, seed: int):
    """Every segment gets `extra` more random ACGT bases, split at random between its two ends (e.g. the 22-base bench
    segments -> 40 bases: what `extract_lr_bc --num-bp-after 38` would cut)."""
    rng = np.random.default_rng(seed + 611953)
    n = offsets.shape[0] - 1
    lens = np.diff(offsets)
    new_off = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(lens + extra, out=new_off[1:])
    out = ACGT[rng.integers(0, 4, size=int(new_off[-1]))]
    left = rng.integers(0, extra + 1, size=n)
    rid = np.repeat(np.arange(n), lens)
    pos = np.arange(seg_bytes.size) - offsets[rid]
    out[new_off[rid] + left[rid] + pos] = seg_bytes
    return out, new_off


def sprinkle_n(seg_bytes: np.ndarray, offsets: np.ndarray, frac: float, seed: int):
    """A copy in which a fraction `frac` of the (non-empty) segments has one base replaced by 'N' (what a basecaller's
    ambiguous call looks like: scored as 'A', and no match may start on the k bytes that cover it — SURVEY.md A.4)."""
    rng = np.random.default_rng(seed + 32452843)
    n = offsets.shape[0] - 1
    lens = np.diff(offsets)
    pick = np.nonzero((rng.random(n) < frac) & (lens > 0))[0]
    out = seg_bytes.copy()
    out[offsets[pick] + (rng.random(pick.size) * lens[pick]).astype(np.int64)] = ord("N")
    return out
