"""SCT_TRACE=1 python tools/trace_probe.py [n]: host-side phase times of sct_match on one page-locked shard (stderr)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from sctagger_b200 import _lib, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_250_000
bc = synth.make_barcodes(25_000, 1002)
sb, off = synth.make_segments_range(bc, 0, n)
psb, poff, out = _lib.Pinned.copy_of(sb), _lib.Pinned.copy_of(off), _lib.ResultBuffers(n)
with _lib.Matcher(bc, 2, 16) as m:
    for _ in range(6):
        t0 = time.perf_counter()
        m.match(psb.array, poff.array, out=out, copy_ties=False)
        print("python-side ms", round(1e3 * (time.perf_counter() - t0), 3), file=sys.stderr)
