"""Executable numpy model of the arithmetic the CUDA kernels perform (sctagger_b200/csrc/match_kernels.cuh).

It exists so that the bit-level design can be checked against the oracle on the CPU box (no GPU there):
tests/test_wm_model.py runs this model against oracle/ on random and golden inputs.  It is NOT a product path.

Recurrence: the k-error Shift-And automaton of Wu & Manber, one register R_d per error level d = 0..k, bit i of
R_d set <=> pattern[0..i] matches some text suffix ending at the current column with <= d edits.  Two barcodes
are interleaved in one 32-bit word (barcode 2w in the even bits, 2w+1 in the odd bits), so "shift by one pattern
position" is `x * 4` and the always-active start state is `+ 3`:  S_d = 4 R_d + 3 v_d  (one IMAD on the GPU).

    R0' = S0 & Eq                                S0' = 4 R0' + 3 v_0[j]
    t   = R_{d-1} | S_{d-1} | S'_{d-1}           (insertion | substitution | deletion)
    Rd' = (S_d & Eq) | t                         Sd' = 4 Rd' + 3 v_d[j]

v_d[j] = [D0[j] <= d] is the start state: D0 is the DP top row of SURVEY.md Appendix A.4 (0 at a valid start,
else D0[j-1]+1); for clean ACGT segments v == 1 everywhere (the `+3` is then an immediate).
A barcode matches with <= d edits iff bit (bl-1) of its field is ever set in R_d.
"""
import numpy as np

M32 = np.uint32(0xFFFFFFFF)


def map_char(b):
    """reference scTagger.py:512-516 — every non-ACGT byte scores as 'A' (0)."""
    out = np.zeros(b.shape, dtype=np.uint8)
    out[b == ord("C")] = 1
    out[b == ord("G")] = 2
    out[b == ord("T")] = 3
    return out


def rc_bytes(b):
    """reference scTagger.py:148-156 — reverse complement, identity on non-ACGT; rows of a (B, bl) matrix."""
    t = np.arange(256, dtype=np.uint8)
    t[ord("A")], t[ord("C")], t[ord("G")], t[ord("T")] = ord("T"), ord("G"), ord("C"), ord("A")
    return t[b[..., ::-1]]


def eq_table(barcodes):
    """(4, W) uint32: interleaved match masks of barcodes 2w (even bits) and 2w+1 (odd bits) per text code."""
    B, bl = barcodes.shape
    W = (B + 1) // 2
    codes = map_char(barcodes)
    tab = np.zeros((4, W), dtype=np.uint32)
    for c in range(4):
        m = (codes == c)
        for i in range(bl):
            bits = m[:, i].astype(np.uint32)
            tab[c, : (B + 1) // 2] |= bits[0::2] << np.uint32(2 * i)
            odd = bits[1::2]
            tab[c, : odd.shape[0]] |= odd << np.uint32(2 * i + 1)
    return tab


def top_row(seg, kpre):
    """D0[0..n] of Appendix A.4 (start j valid iff j+kpre <= n and seg[j:j+kpre] all upper-case ACGT)."""
    n = len(seg)
    INF = 1 << 20
    acgt = np.isin(seg, np.frombuffer(b"ACGT", dtype=np.uint8))
    d0 = np.zeros(n + 1, dtype=np.int64)
    prev = INF
    for j in range(n + 1):
        valid = j + kpre <= n and bool(acgt[j:j + kpre].all())
        prev = 0 if valid else min(INF, prev + 1)
        d0[j] = prev
    return d0


def wm_levels(tab, codes, d0, K, bl):
    """Run the automaton for all words.  Returns acc (K+1, W) uint32: OR over columns of R_d.
    codes: text codes per column; d0: top row (len n+1) or None for the clean case (all zeros)."""
    W = tab.shape[1]
    n = len(codes)
    if d0 is None:
        d0 = np.zeros(n + 1, dtype=np.int64)
    R, S = [], []
    for d in range(K + 1):
        # bit i of level d initially set iff D0[0] + i + 1 <= d
        nb = max(0, min(16, d - int(d0[0]))) if d0[0] <= d else 0
        r = np.uint32((1 << (2 * nb)) - 1)
        R.append(np.full(W, r, dtype=np.uint32))
        S.append((R[d] * np.uint32(4) + np.uint32(3 if d0[0] <= d else 0)) & M32)
    acc = [np.zeros(W, dtype=np.uint32) for _ in range(K + 1)]
    for j in range(n):
        eq = tab[codes[j]]
        v = [np.uint32(3 if d0[j + 1] <= d else 0) for d in range(K + 1)]
        R0n = S[0] & eq
        S0n = (R0n * np.uint32(4) + v[0]) & M32
        newR, newS = [R0n], [S0n]
        for d in range(1, K + 1):
            t = R[d - 1] | S[d - 1] | newS[d - 1]
            Rn = (S[d] & eq) | t
            newR.append(Rn)
            newS.append((Rn * np.uint32(4) + v[d]) & M32)
        R, S = newR, newS
        for d in range(K + 1):
            acc[d] |= R[d]
    return np.stack(acc)


def levels_to_dist(acc, bl, B, K):
    """per barcode: smallest d whose level has bit (bl-1) of its field set, K+1 if none."""
    W = acc.shape[1]
    out = np.full(2 * W, K + 1, dtype=np.int64)
    for d in range(K, -1, -1):
        even = (acc[d] >> np.uint32(2 * (bl - 1))) & np.uint32(1)
        odd = (acc[d] >> np.uint32(2 * (bl - 1) + 1)) & np.uint32(1)
        out[0::2][even == 1] = d
        out[1::2][odd == 1] = d
    return out[:B]


def match_model(barcodes, seg_bytes, offsets, max_error, threads, fast_path=True):
    """Model of the whole match stage.  fast_path=True mirrors the kernel split: clean segments of <= 32 bases use
    the forward table against T and reverse(T) with complemented codes (RC strand); everything else uses the
    explicit forward + reverse-complement tables with the start-state row."""
    B, bl = barcodes.shape
    K = max_error
    kpre, p = 0, 1
    while threads > p:
        p *= 4
        kpre += 1
    tabF = eq_table(barcodes)
    tabR = eq_table(rc_bytes(barcodes))
    clean_bc = bool(np.isin(barcodes, np.frombuffer(b"ACGT", dtype=np.uint8)).all())
    n = len(offsets) - 1
    min_ed = np.full(n, -1, dtype=np.int32)
    n_ties = np.zeros(n, dtype=np.int32)
    lists = []
    for r in range(n):
        seg = seg_bytes[offsets[r]:offsets[r + 1]]
        L = len(seg)
        if L < bl - K:
            lists.append(np.zeros(0, dtype=np.int32))
            continue
        codes = map_char(seg)
        clean = bool(np.isin(seg, np.frombuffer(b"ACGT", dtype=np.uint8)).all())
        if fast_path and clean and clean_bc and L <= 32:
            accF = wm_levels(tabF, codes, None, K, bl)
            accR = wm_levels(tabF, (3 - codes)[::-1], None, K, bl)
        else:
            d0 = top_row(seg, kpre)
            accF = wm_levels(tabF, codes, d0, K, bl)
            accR = wm_levels(tabR, codes, d0, K, bl)
        dF = levels_to_dist(accF, bl, B, K)
        dR = levels_to_dist(accR, bl, B, K)
        best = int(min(dF.min(), dR.min()))
        if best > K:
            lists.append(np.zeros(0, dtype=np.int32))
            continue
        ids = np.concatenate([(np.nonzero(dR == best)[0] << 1), (np.nonzero(dF == best)[0] << 1) | 1])
        ids.sort()
        min_ed[r] = best
        n_ties[r] = len(ids)
        lists.append(ids.astype(np.int32))
    tie_off = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(n_ties, out=tie_off[1:])
    tie_ids = np.concatenate(lists) if lists else np.zeros(0, dtype=np.int32)
    return min_ed, n_ties, tie_off, tie_ids.astype(np.int32)
