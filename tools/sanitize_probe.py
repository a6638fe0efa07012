"""Small end-to-end exercise of every kernel family for compute-sanitizer (run on the GPU box):
    compute-sanitizer --tool memcheck python tools/sanitize_probe.py      (closed on the round-2 pool: run plain there)
    compute-sanitizer --tool racecheck python tools/sanitize_probe.py
Results are compared with the oracle so that a silent corruption shows up too."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import util
from oracle import oracle
from sctagger_b200 import _lib, synth

rng = np.random.default_rng(5)
for mr, t in ((2, 16), (3, 64), (1, 4)):
    bc, sb, off = util.random_case(rng, 1500, 1200, bl=16, dirty=0.1, lower=True, max_len=31)
    bc2, sb2, off2 = util.long_case(rng, 1500, 300, bl=16, min_len=20, max_len=120, dirty=0.3)
    sb = np.concatenate([sb, sb2]); off = np.concatenate([off, off[-1] + off2[1:]])
    want = oracle.match(bc, sb, off, mr, threads=t, algo="dp")
    for opts in ({}, {"seed_deepening": 0, "seed_prefilter": 1, "seed_direct": 0}, {"path": 3}, {"chunk_segments": 257}):
        got = _lib.match_segments(bc, sb, off, max_error=mr, threads=t, **opts)
        assert np.array_equal(got.min_ed, want[0]) and np.array_equal(got.n_ties, want[1]) and np.array_equal(got.tie_ids, want[3]), (mr, opts)
rb, off = synth.make_reads(600, 7)
long_read = synth.ACGT[rng.integers(0, 4, size=40_000)]
rb = np.concatenate([rb, long_read]); off = np.append(off, off[-1] + 40_000)
want = oracle.align(synth.ADAPTER, rb, off, algo="bv")
got = _lib.align_reads(synth.ADAPTER, rb, off)
assert all(np.array_equal(x, y) for x, y in zip(got, want))
s, d = _lib.align_strand_dist_noalloc(synth.ADAPTER, rb, off[:601])
assert np.array_equal(s, want[0][:600]) and np.array_equal(d, want[1][:600])
print("sanitize probe ok")
