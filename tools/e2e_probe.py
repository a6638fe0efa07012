"""Where the end-to-end time of Matcher.match goes (run on the GPU box with SCT_TRACE=1)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sctagger_b200 import _lib, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
opts = dict(kv.split("=") for kv in sys.argv[2:])
bc = synth.make_barcodes(25_000, 1002)
sb, off = synth.make_segments(bc, n, 2002)
with _lib.Matcher(bc, 2, 16, **{k: int(v) for k, v in opts.items()}) as m:
    for i in range(5):
        t0 = time.perf_counter()
        r = m.match(sb, off)
        print(f"match() {1e3 * (time.perf_counter() - t0):.3f} ms", flush=True)
