"""numpy / pure-Python twin of the *seeded* match kernel (sctagger_b200/csrc/seed_kernels.cuh), bit for bit:

  1. pigeonhole seeds: the barcode is cut into K+1 disjoint pieces; a barcode within K edits of some substring of the
     segment has at least one piece occurring *exactly* in the segment, at a text position j with
     o_p - K <= j <= n - (bl - o_p) + K (o_p = offset of the piece in the barcode).  Pieces are indexed per piece in
     bucket tables (bucket = 2-bit code of the piece), so the candidates of a segment are a handful of bucket ranges.
  2. verification: every candidate (barcode, strand) is scored exactly against the whole segment by a k-error
     Shift-And automaton with the ROLES SWAPPED — the state vector runs over the segment's positions (one 32-bit
     register, n <= 31) and one barcode character is consumed per step:
         R0' = (R0 << 1) & Eq[c]                       Eq[c] bit j = [T[j-1] == c]
         Rd' = ((Rd << 1) & Eq[c]) | R_{d-1} | (R_{d-1} << 1) | (R'_{d-1} << 1)
     start: R_d = all positions 0..n (a match may start anywhere); after bl steps the barcode matches within d edits
     iff R_d & mask(n+1) != 0.
  Reverse strand: d(revcomp(P), T) = d(P, revcomp(T)), so revcomp(T) is scored against the forward barcode.

Only clean-ACGT segments of <= 31 bases take this path in the library; the model mirrors exactly that and returns
None for other segments.  Test infrastructure (CPU): checked against the oracle in tests/test_index_model.py.
"""
import numpy as np

CODE = {65: 0, 67: 1, 71: 2, 84: 3}


def piece_layout(bl, K, max_piece=8):
    """K+1 disjoint pieces covering the barcode left to right (lengths differ by at most 1, capped at max_piece).
    -> list of (offset, length).  Mirrors sct::make_pieces in csrc/sctagger_b200.cu."""
    P = K + 1
    base, extra = divmod(bl, P)
    out, o = [], 0
    for p in range(P):
        ln = base + (1 if p >= P - extra else 0)      # longer pieces last
        out.append((o, min(ln, max_piece)))
        o += ln
    return out


def pack(codes):
    v = 0
    for i, c in enumerate(codes):
        v |= int(c) << (2 * i)
    return v


def build_index(bc_codes, K):
    """bc_codes: (B, bl) array of 2-bit codes.  -> pieces, per piece dict bucket -> list of barcode ids (ascending)."""
    B, bl = bc_codes.shape
    pieces = piece_layout(bl, K)
    tables = []
    for (o, ln) in pieces:
        t = {}
        for b in range(B):
            t.setdefault(pack(bc_codes[b, o:o + ln]), []).append(b)
        tables.append(t)
    return pieces, tables


def verify(word_codes, text_codes, K):
    """Exact semi-global distance of the barcode (codes) inside the text (codes), capped: -> level 0..K or K+1."""
    n = len(text_codes)
    full = (1 << (n + 1)) - 1
    eq = [0, 0, 0, 0]
    for j, c in enumerate(text_codes):
        eq[c] |= 1 << (j + 1)
    M32 = 0xFFFFFFFF
    R = [full] * (K + 1)
    for c in word_codes:
        e = eq[int(c)]
        S = [(r << 1) & M32 for r in R]
        nR = [S[0] & e]
        for d in range(1, K + 1):
            nR.append(((S[d] & e) | R[d - 1] | S[d - 1] | ((nR[d - 1] << 1) & M32)) & M32)
        R = nR
    for d in range(K + 1):
        if R[d] & full:
            return d
    return K + 1


def match_segment(bc_codes, index, seg, K, stats=None):
    """-> (min_ed or -1, sorted unique tie ids) for one clean segment, or None if the segment is not on this path."""
    pieces, tables = index
    bl = bc_codes.shape[1]
    n = len(seg)
    if n > 31 or any(b not in CODE for b in seg):
        return None
    if n < bl - K:
        return -1, []
    fwd = [CODE[b] for b in seg]
    rc = [3 - c for c in reversed(fwd)]
    best, ties = K + 1, set()
    for strand, text in ((1, fwd), (0, rc)):
        for (o, ln), table in zip(pieces, tables):
            lo, hi = max(0, o - K), min(n - ln, n - (bl - o) + K)
            for j in range(lo, hi + 1):
                for b in table.get(pack(text[j:j + ln]), ()):
                    if stats is not None:
                        stats["candidates"] = stats.get("candidates", 0) + 1
                    lv = verify(bc_codes[b], text, K)
                    if lv < best:
                        best, ties = lv, set()
                    if lv == best and lv <= K:
                        ties.add((b << 1) | strand)
    if best > K:
        return -1, []
    return best, sorted(ties)
