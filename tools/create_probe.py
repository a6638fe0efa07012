"""Where the first call of a fresh handle spends its time: context, sct_create, first / second sct_match on pageable arrays."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from sctagger_b200 import _lib, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
bc = synth.make_barcodes(25_000, 1002)
sb, off = synth.make_segments_range(bc, 0, n)
t0 = time.perf_counter(); _lib.device_count(); m0 = _lib.Matcher(bc, 2, 16); t1 = time.perf_counter()
print("first create (CUDA context + indexes) s", round(t1 - t0, 3))
for i in range(3):
    t0 = time.perf_counter(); r = m0.match(sb, off); print("match", i, "s", round(time.perf_counter() - t0, 3))
m0.close()
for i in range(2):
    t0 = time.perf_counter(); m = _lib.Matcher(bc, 2, 16); t1 = time.perf_counter(); r = m.match(sb, off); t2 = time.perf_counter(); m.close(); t3 = time.perf_counter()
    print("create s", round(t1 - t0, 3), "first match s", round(t2 - t1, 3), "close s", round(t3 - t2, 3))
