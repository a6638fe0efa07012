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


def start_levels(seg, kpre):
    """D0 of SURVEY.md Appendix A.4 for positions 0..n: 0 where a match may start (the next kpre bytes exist and are all
    upper-case ACGT — the reference's -t prefix rule), else the previous value + 1.  Mirrors prep_seed_kernel."""
    n = len(seg)
    out, prev = [], 1000
    for j in range(n + 1):
        valid = j + kpre <= n and all(b in CODE for b in seg[j:j + kpre])
        prev = 0 if valid else min(prev + 1, 1000)
        out.append(prev)
    return out


def verify_restricted(word_codes, text_codes, K, init):
    """verify() with per-level start states: init[d] bit j = a match may start at text position j with d edits spent."""
    n = len(text_codes)
    full = (1 << (n + 1)) - 1
    eq = [0, 0, 0, 0]
    for j, c in enumerate(text_codes):
        eq[c] |= 1 << (j + 1)
    M32 = 0xFFFFFFFF
    R = list(init[:K + 1])
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


def match_segment_windowed(bc_codes, index, seg, K, kpre):
    """Twin of match_seed_kernel<GEN = true>: ANY segment.  Non-ACGT bytes are scored as A; overlapping windows of <= 31
    bases, window step 32 - bl - K; per window the pigeonhole seeds of both strands (restriction ignored: necessary
    condition) and a verification on the FORWARD text with the window's slice of the start levels — the reverse strand as
    revcomp(barcode).  -> (min_ed or -1, sorted unique tie ids)"""
    pieces, tables = index
    bl = bc_codes.shape[1]
    n_total = len(seg)
    if n_total < bl - K:
        return -1, []
    codes = [CODE.get(b, 0) for b in seg]
    d0 = start_levels(seg, kpre)
    stride = 32 - bl - K
    best, ties = K + 1, set()
    ws = 0
    while True:
        n = min(31, n_total - ws)
        fwd = codes[ws:ws + n]
        rc = [3 - c for c in reversed(fwd)]
        init = [sum(1 << j for j in range(n + 1) if d0[ws + j] <= d) for d in range(4)]
        for strand, text in ((0, rc), (1, fwd)):
            for (o, ln), table in zip(pieces, tables):
                lo, hi = max(0, o - K), min(n - ln, n - (bl - o) + K)
                for j in range(lo, hi + 1):
                    for b in table.get(pack(text[j:j + ln]), ()):
                        word = bc_codes[b] if strand else [3 - int(c) for c in reversed(bc_codes[b])]
                        lv = verify_restricted(word, fwd, K, init)
                        if lv < best:
                            best, ties = lv, set()
                        if lv == best and lv <= K:
                            ties.add((b << 1) | strand)
        if ws + 31 >= n_total:
            break
        ws += stride
    if best > K:
        return -1, []
    return best, sorted(ties)


# ---- responsible-seed pre-check (seed_precheck in csrc/seed_kernels.cuh, rules set in sct_create) ------------------------
def precheck_rule(bl, K, q):
    """-> (go_left, characters read, level tested) for piece q of the error-bound-K index (bl = 16 only), or None."""
    if bl != 16 or K < 1 or K > 3:
        return None
    if K == 1:
        return (q == 1, 6, 1)
    if K == 2:
        return (q >= 1, 6 if q == 2 else 5, 2 if q == 2 else 1)
    return (q % 2 == 1, 4, 1)


def anchored_cost(pat, text):
    """min over prefixes text[:L] of the edit distance between pat and text[:L] (anchored start, free end in the text)."""
    prev = list(range(len(text) + 1))                  # pat[:0] against text[:L]: L insertions
    for i, c in enumerate(pat, 1):
        cur = [i] + [0] * len(text)
        for L in range(1, len(text) + 1):
            cur[L] = min(prev[L - 1] + (c != text[L - 1]), prev[L] + 1, cur[L - 1] + 1)
        prev = cur
    return min(prev)


def precheck_passes(bc_row, text, bl, K, q, piece, j):
    """Twin of seed_precheck for one candidate: piece q = (offset, length) of barcode bc_row occurs exactly at text[j:]."""
    rule = precheck_rule(bl, K, q)
    if rule is None:
        return True
    go_left, need, level = rule
    o, ln = piece
    if go_left:
        pat = [int(c) for c in bc_row[o - need:o]][::-1]           # the barcode characters before the piece, outwards
        side = list(text[:j])[::-1]
    else:
        pat = [int(c) for c in bc_row[o + ln:o + ln + need]]
        side = list(text[j + ln:])
    return anchored_cost(pat, side) <= level
