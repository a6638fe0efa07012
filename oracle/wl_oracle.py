"""TEST INFRASTRUCTURE ONLY — CPU restatement of the whitelist search of `extract_sr_bc_from_lr`
(reference scTagger.py:792-844).  Never imported by the product (sctagger_b200/).

The arithmetic of this stage lives in pyahocorasick (third-party, unpinned in env.yaml, absent from /root/reference and
not installable here), so **parity with pyahocorasick is unpinned**; what is restated are its documented semantics:
  * `A.add_word(key, value)` on an existing key replaces the value           (forward calls scTagger.py:804-805, then the
    reverse-complement calls scTagger.py:807-808, so an RC string that equals a forward string wins);
  * `A.iter(text)` yields `(end_index, value)` for every occurrence of every key, ordered by end index.
`Automaton` below is a drop-in for the three methods the reference calls; tools/make_golden_wl.py runs the UNMODIFIED
reference with it, which pins everything of the stage except pyahocorasick itself.
"""
from collections import Counter

import numpy as np


class Automaton:
    """The subset of ahocorasick.Automaton the reference uses (scTagger.py:802-809, 823), by plain dictionary lookup."""

    def __init__(self):
        self.words = {}

    def add_word(self, key, value):
        if key == "":
            return False
        new = key not in self.words
        self.words[key] = value
        return new

    def make_automaton(self):
        self.lengths = sorted({len(k) for k in self.words})

    def iter(self, text):
        for end in range(len(text)):
            for ln in self.lengths:               # same end index: shorter keys first (order is irrelevant to the stage)
                start = end - ln + 1
                if start >= 0:
                    v = self.words.get(text[start:end + 1])
                    if v is not None:
                        yield end, v


def rev_compl(s):
    """reference scTagger.py:148-156 (identity on bytes outside ACGT)."""
    m = {"A": "T", "C": "G", "G": "C", "T": "A"}
    return "".join(m.get(c, c) for c in reversed(s))


def search(barcodes, segments):
    """scTagger.py:802-827 -> (n_hits per segment, abs(value) of the first hit or -1, Counter in insertion order)."""
    A = Automaton()
    for idx, bc in enumerate(barcodes):
        A.add_word(bc, idx)
    for idx, bc in enumerate(barcodes):
        A.add_word(rev_compl(bc), -idx)
    A.make_automaton()
    C = Counter()
    n_hits = np.zeros(len(segments), dtype=np.int32)
    first = np.full(len(segments), -1, dtype=np.int32)
    for r, seg in enumerate(segments):
        hits = tuple(A.iter(seg))
        n_hits[r] = len(hits)
        if hits:
            first[r] = abs(hits[0][1])
        if len(hits) > 1:
            continue
        for _, bc in hits:
            C[abs(bc)] += 1
    return n_hits, first, C


def select_top(C, thresh, step_size, max_barcode_cnt):
    """scTagger.py:829-836: stable sort by count (descending), cap, then cut at the first step below `thresh`."""
    sorted_bc = sorted(C.items(), reverse=True, key=lambda x: x[1])[:max_barcode_cnt]
    total = sum(c for _, c in sorted_bc)
    for last_idx in range(0, len(sorted_bc), step_size):
        percent = sum(c for _, c in sorted_bc[last_idx:last_idx + step_size]) / total
        if percent < thresh:
            break
    return sorted_bc[:last_idx + step_size]          # UnboundLocalError when nothing was counted, like the reference


def search_numpy(barcodes, seg_bytes, offsets):
    """Vectorised twin for large inputs (one word length, ACGT-only whitelist): -> (n_hits, first) per segment."""
    L = len(barcodes[0])
    code = np.full(256, -1, dtype=np.int64)
    for i, c in enumerate(b"ACGT"):
        code[c] = i
    table = {}
    for idx, bc in enumerate(barcodes):
        table[bc] = idx
    for idx, bc in enumerate(barcodes):
        table[rev_compl(bc)] = idx
    keys = np.array([sum(int(code[ord(ch)]) << (2 * (L - 1 - i)) for i, ch in enumerate(k)) for k in table], dtype=np.int64)
    vals = np.array(list(table.values()), dtype=np.int64)
    order = np.argsort(keys)
    keys, vals = keys[order], vals[order]
    sb = np.asarray(seg_bytes, dtype=np.uint8)
    off = np.asarray(offsets, dtype=np.int64)
    n = off.shape[0] - 1
    c = code[sb]
    tot = sb.shape[0]
    n_hits = np.zeros(n, dtype=np.int32)
    first = np.full(n, -1, dtype=np.int32)
    if tot < L:
        return n_hits, first
    w = np.zeros(tot - L + 1, dtype=np.int64)
    ok = np.ones(tot - L + 1, dtype=bool)
    for i in range(L):
        ci = c[i:tot - L + 1 + i]
        ok &= ci >= 0
        w = (w << 2) | np.maximum(ci, 0)
    seg_of = np.searchsorted(off, np.arange(tot - L + 1), side="right") - 1
    ok &= (np.arange(tot - L + 1) + L) <= off[np.minimum(seg_of + 1, n)]
    pos = np.searchsorted(keys, w)
    pos = np.minimum(pos, keys.shape[0] - 1)
    hit = ok & (keys[pos] == w)
    hs = seg_of[hit]
    np.add.at(n_hits, hs, 1)
    hv = vals[pos[hit]]
    uniq, first_pos = np.unique(hs, return_index=True)
    first[uniq] = hv[first_pos]
    return n_hits, first
