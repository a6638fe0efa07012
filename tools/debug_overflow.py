import sys, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import util
from oracle import oracle
from sctagger_b200 import _lib
rng = np.random.default_rng(5)
one = np.frombuffer(b"ACGTTGCAAGGCTTAC", dtype=np.uint8)
for ncopy in (100, 230, 700):
    bc = np.tile(one, (ncopy, 1))
    other = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, size=(100, 16))]
    bc = np.concatenate([bc[:ncopy // 2], other, bc[ncopy // 2:]])
    segs = [b"GG" + one.tobytes() + b"TTA", b"ACGTACGTACGTACGTACGTAA", one.tobytes()[:15] + b"CCCCC"] * 40
    off = np.zeros(len(segs) + 1, np.int64)
    np.cumsum([len(s) for s in segs], out=off[1:])
    sb = np.frombuffer(b"".join(segs), np.uint8)
    want = util.result_from_tuple(oracle.match(bc, sb, off, 2, threads=16, algo="dp"))
    for opts in ({}, {"tie_capacity_per_segment": 1}, {"path": 1}, {"path": 2}):
        with _lib.Matcher(bc, 2, 16, **opts) as m:
            got = m.match(sb, off)
            print(ncopy, opts, "equal", got.equals(want), got.min_ed[:6], got.n_ties[:6], want.n_ties[:6], m.stats())
