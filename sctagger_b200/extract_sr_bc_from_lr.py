"""Host-side mirror of the reference's `extract_sr_bc_from_lr` stage (reference scTagger.py:792-844): the exact search of
every whitelist barcode (and its reverse complement) in the long-read segments written by `extract_lr_bc`, the count of
segments that hold exactly one occurrence, and the selection of the top barcodes — the `-sr` input of `match_trie` when
no short reads exist (Snakefile:91-119).

    read_whitelist(path)                 <- scTagger.py:793-799
    read_segments(path)                  <- scTagger.py:811-822   (rows with location 'NA' are skipped)
    count_barcodes(...)                  <- scTagger.py:802-827   (the seam: here it calls the CUDA library)
    select_top(...)                      <- scTagger.py:829-836
    extract_sr_bc_from_lr(args)          <- scTagger.py:792-844

The Aho-Corasick automaton is replaced by a hash table in HBM probed once per window (sct_wl_* in
include/sctagger_b200.h); there is no CPU fallback.
"""
from __future__ import annotations

import sys

import numpy as np

from .match_trie import GZIP_LEVEL


def read_whitelist(path: str, threads=None):
    """Every line minus its last character (`l[:-1]`, scTagger.py:798) -> (word_bytes, word_offsets)."""
    from . import _lib
    print(f"Reading whiltelist barcodes from: {path}")
    return _lib.io_read_whitelist(path, threads)


def read_segments(path: str, threads=None):
    """`_,_,p,seg = l.rstrip('\\n').split('\\t')`; rows with p == 'NA' are skipped (scTagger.py:819-822).
    A row that does not have exactly four columns raises ValueError like the reference."""
    from . import _lib
    return _lib.io_read_lr_bc_segments(path, threads)


def count_barcodes(word_bytes, word_offsets, seg_bytes, seg_offsets, threads: int = 16, device: int = 0):
    """The Counter `C` of scTagger.py:818-827 in insertion order -> (barcode index int32[u], count int64[u])."""
    from . import _lib
    with _lib.Whitelist(word_bytes, word_offsets, threads=threads, device=device) as wl:
        wl.count(seg_bytes, seg_offsets)
        return wl.result()


def select_top(idx: np.ndarray, cnt: np.ndarray, thresh: float, step_size: int, max_barcode_cnt: int):
    """scTagger.py:829-836: stable sort by count descending (ties keep Counter order), cap at --max-barcode-cnt, then keep
    whole steps of --step-size until a step carries less than --thresh of the total."""
    order = np.argsort(-cnt, kind="stable")[:max_barcode_cnt]
    idx, cnt = idx[order], cnt[order]
    total = int(cnt.sum())
    if idx.shape[0] == 0:
        # the reference leaves `last_idx` unbound when nothing was counted (scTagger.py:832-836)
        raise UnboundLocalError("cannot access local variable 'last_idx' where it is not associated with a value")
    last_idx = 0
    for last_idx in range(0, idx.shape[0], step_size):
        percent = int(cnt[last_idx:last_idx + step_size].sum()) / total
        if percent < thresh:
            break
    return idx[:last_idx + step_size], cnt[:last_idx + step_size]


def format_rows(word_bytes, word_offsets, idx, cnt) -> bytes:
    """`barcode \\t count` per selected barcode (scTagger.py:843-844)."""
    raw = np.asarray(word_bytes, dtype=np.uint8).tobytes()
    o = np.asarray(word_offsets)
    return b"".join(b"%s\t%d\n" % (raw[o[b]:o[b + 1]], c) for b, c in zip(idx.tolist(), cnt.tolist()))


def extract_sr_bc_from_lr_bytes(args) -> bytes:
    threads = getattr(args, "threads", None) or 16
    wb, wo = read_whitelist(args.barcode_whitelist, threads)
    print(f"\n=====\nMatching exact barcodes on long-reads: {args.input}")
    sb, so = read_segments(args.input, threads)
    devices = getattr(args, "devices", None) or [0]
    idx, cnt = count_barcodes(wb, wo, sb, so, threads=threads, device=devices[0])
    print(f"\n=====\nFound {idx.shape[0]:,} unique barcodes on long-reads")
    idx, cnt = select_top(idx, cnt, args.thresh, args.step_size, args.max_barcode_cnt)
    print(f"\n=====\nWriting the top {idx.shape[0]} barcodes")
    return format_rows(wb, wo, idx, cnt)


def extract_sr_bc_from_lr(args):
    """Stage driver, same steps as reference scTagger.py:792-844; gzip whenever -o is given (scTagger.py:839-840)."""
    import gzip
    data = extract_sr_bc_from_lr_bytes(args)
    if args.outfile:
        with gzip.open(args.outfile, "wb", compresslevel=GZIP_LEVEL) as f:
            f.write(data)
    else:
        sys.stdout.buffer.write(data)
        sys.stdout.buffer.flush()
