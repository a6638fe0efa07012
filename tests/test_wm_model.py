"""The bit-level model of the kernels' recurrence (tests/wm_model.py) against the oracle — CPU only."""
import numpy as np
import pytest

import util
import wm_model
from oracle import oracle


@pytest.mark.parametrize("seed", range(10))
def test_model_matches_oracle_random(seed):
    rng = np.random.default_rng(500 + seed)
    bl = [16, 16, 16, 12, 8, 16, 15, 16, 10, 16][seed]
    mr = [2, 3, 0, 1, 2, 1, 3, 2, 3, 2][seed]
    t = [16, 1, 4, 17, 16, 5, 16, 64, 1, 16][seed]
    dirty = [0.0, 0.3, 0.3, 0.2, 0.2, 0.0, 0.3, 0.5, 0.2, 0.0][seed]
    bc, sb, off = util.random_case(rng, 37, 120, bl=bl, dirty=dirty, lower=True, max_len=[40, 40, 40, 40, 40, 70, 40, 40, 40, 30][seed])
    if seed in (3, 7):   # dirty barcodes
        bc = bc.copy()
        bc[5, 2] = ord("N")
        bc[11, bl - 1] = ord("n")
    want = oracle.match(bc, sb, off, mr, threads=t, algo="dp")
    got = wm_model.match_model(bc, sb, off, mr, t)
    for g, w in zip(got, want):
        assert np.array_equal(g, w)
    # generic path alone must agree too
    got2 = wm_model.match_model(bc, sb, off, mr, t, fast_path=False)
    for g, w in zip(got2, want):
        assert np.array_equal(g, w)


@pytest.mark.parametrize("c", [c for c in util.golden_cases() if c["dir"] in ("edge16", "edge8", "ties16") and c["t"] in (1, 16)],
                         ids=util.case_id)
def test_model_matches_reference_golden(c):
    from sctagger_b200 import match_trie as mt
    barcodes, reads, expected = util.load_case(c)
    bm = mt.barcode_matrix(barcodes, c["bl"])
    res = util.result_from_tuple(wm_model.match_model(bm, reads.seg_bytes, reads.offsets, c["mr"], c["t"]))
    assert mt.format_rows(reads.names, reads.seg_bytes, reads.offsets, res, barcodes) == expected
