"""CPU: the numpy/Python twin of the seeded kernel (pigeonhole seeds + role-swapped Shift-And verification) against
the oracle, so that the arithmetic is proven before any GPU time is spent."""
import numpy as np
import pytest

import index_model as im
import util
from oracle import oracle


@pytest.mark.parametrize("bl,K,t", [(16, 2, 16), (16, 3, 1), (16, 1, 4), (16, 0, 16), (12, 2, 16), (10, 3, 1), (8, 1, 16), (15, 2, 5)])
def test_model_equals_oracle(bl, K, t):
    rng = np.random.default_rng(bl * 100 + K)
    bc, sb, off = util.random_case(rng, 120, 400, bl=bl, dirty=0.0, max_len=31)
    want = util.result_from_tuple(oracle.match(bc, sb, off, K, threads=t, algo="dp"))
    codes = np.vectorize(im.CODE.get)(bc).astype(np.uint8)
    index = im.build_index(codes, K)
    raw = sb.tobytes()
    n_checked = 0
    for r in range(len(off) - 1):
        got = im.match_segment(codes, index, raw[off[r]:off[r + 1]], K)
        if got is None:
            continue
        n_checked += 1
        assert got[0] == want.min_ed[r], r
        assert got[1] == want.ties(r).tolist(), r
    assert n_checked > 300


def test_piece_layout_is_disjoint_and_has_k_plus_1_pieces():
    for bl in range(4, 17):
        for K in range(0, min(4, bl)):
            p = im.piece_layout(bl, K)
            assert len(p) == K + 1
            end = 0
            for o, ln in p:
                assert o >= end and 1 <= ln <= 8
                end = o + ln
            assert end <= bl


@pytest.mark.parametrize("bl,K,t", [(16, 2, 16), (16, 3, 64), (16, 1, 4), (12, 2, 17), (14, 3, 1), (16, 2, 5)])
def test_windowed_model_equals_oracle_on_long_and_dirty_segments(bl, K, t):
    """The windowed variant (any length, non-ACGT bytes, start restrictions of the -t prefix rule) against the oracle."""
    rng = np.random.default_rng(bl * 1000 + K * 10 + t)
    bc, sb, off = util.long_case(rng, 150, 260, bl=bl, min_len=bl - K - 1, max_len=110, dirty=0.5)
    want = util.result_from_tuple(oracle.match(bc, sb, off, K, threads=t, algo="dp"))
    kpre = 0
    while 4 ** kpre < t:
        kpre += 1
    codes = np.vectorize(im.CODE.get)(bc).astype(np.uint8)
    index = im.build_index(codes, K)
    raw = sb.tobytes()
    for r in range(len(off) - 1):
        got = im.match_segment_windowed(codes, index, raw[off[r]:off[r + 1]], K, kpre)
        assert got[0] == want.min_ed[r], r
        assert got[1] == want.ties(r).tolist(), r


@pytest.mark.parametrize("K,seed", [(1, 1), (2, 2), (3, 3), (2, 4), (3, 5)])
def test_responsible_seed_precheck_never_loses_a_match(K, seed):
    """The anchored pre-checks (K = 1: sibling half; K = 2: priority piece 2 > 1 > 0 with the 5-mers testing their
    neighbour within one edit; K = 3: sibling 4-mer within one edit) are necessary conditions only for the RESPONSIBLE seed
    of an alignment: every (barcode, strand) within K edits of a segment must still be reached through at least one of
    its seed occurrences that passes.  Checked here in plain Python, independent of the CUDA code."""
    rng = np.random.default_rng(8800 + seed)
    bc, sb, off = util.random_case(rng, 150, 500, bl=16, dirty=0.0, max_len=31)
    codes = np.vectorize(im.CODE.get)(bc).astype(np.uint8)
    pieces, tables = im.build_index(codes, K)
    raw = sb.tobytes()
    n_true = n_cand = n_pass = 0
    for r in range(len(off) - 1):
        seg = raw[off[r]:off[r + 1]]
        n = len(seg)
        if n < 16 - K:
            continue
        fwd = [im.CODE[b] for b in seg]
        for strand, text in ((1, fwd), (0, [3 - c for c in reversed(fwd)])):
            true, reached = set(), set()
            for q, ((o, ln), table) in enumerate(zip(pieces, tables)):
                lo, hi = max(0, o - K), min(n - ln, n - (16 - o) + K)
                for j in range(lo, hi + 1):
                    for b in table.get(im.pack(text[j:j + ln]), ()):
                        n_cand += 1
                        if im.verify(codes[b], text, K) <= K:
                            true.add(b)
                        if im.precheck_passes(codes[b], text, 16, K, q, (o, ln), j):
                            n_pass += 1
                            reached.add(b)
            assert true <= reached, (r, strand, sorted(true - reached))
            n_true += len(true)
    assert n_true > 100 and n_pass < n_cand          # many real matches were checked, and the filter does filter


def test_anchored_verification_of_error_bound_1_is_exact():
    """Phase 1 (pieces 8 + 8) scores a candidate ANCHORED at its seed: the other half, 8 characters outwards from the seed
    boundary, within 0 or 1 edits (seed_anchor_level1).  Through any seed occurrence that can only over-estimate the
    barcode's distance, and through the responsible one it is exact: the minimum over a barcode's seed occurrences equals
    the true distance whenever that is <= 1."""
    rng = np.random.default_rng(4711)
    bc, sb, off = util.random_case(rng, 200, 700, bl=16, dirty=0.0, max_len=31)
    codes = np.vectorize(im.CODE.get)(bc).astype(np.uint8)
    pieces, tables = im.build_index(codes, 1)
    assert pieces == [(0, 8), (8, 8)]
    raw = sb.tobytes()
    n_checked = 0
    for r in range(len(off) - 1):
        seg = raw[off[r]:off[r + 1]]
        n = len(seg)
        if n < 15:
            continue
        fwd = [im.CODE[b] for b in seg]
        for text in (fwd, [3 - c for c in reversed(fwd)]):
            best = {}
            for q, ((o, ln), table) in enumerate(zip(pieces, tables)):
                for j in range(max(0, o - 1), min(n - ln, n - (16 - o) + 1) + 1):
                    for b in table.get(im.pack(text[j:j + ln]), ()):
                        if q == 0:
                            lv = im.anchored_cost([int(c) for c in codes[b][8:]], list(text[j + 8:]))
                        else:
                            lv = im.anchored_cost([int(c) for c in codes[b][:8]][::-1], list(text[:j])[::-1])
                        lv = min(lv, 2)
                        true = im.verify(codes[b], text, 1)
                        assert lv >= true, (r, b, q, j)                      # never an under-estimate
                        best[b] = (min(best.get(b, (2, 2))[0], lv), true)
            for b, (lv, true) in best.items():
                if true <= 1:
                    assert lv == true, (r, b)
                    n_checked += 1
    assert n_checked > 200
