"""Sweep tuning options of the seeded kernel on the bench workload (run on the GPU box):
    python tools/tune_seed.py [--segments N] [--barcodes B] [--max-error K] key=v1,v2,... """
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sctagger_b200 import _lib, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--segments", type=int, default=1_000_000)
ap.add_argument("--barcodes", type=int, default=25_000)
ap.add_argument("--max-error", type=int, default=2)
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("sweep", nargs="*")
a = ap.parse_args()
bc = synth.make_barcodes(a.barcodes, 1002)
sb, off = synth.make_segments(bc, a.segments, 2002)
combos = [{}]
for s in a.sweep:
    k, vs = s.split("=")
    combos = [dict(c, **{k: int(v)}) for c in combos for v in vs.split(",")]
for c in combos:
    with _lib.Matcher(bc, a.max_error, 16, **c) as m:
        m.stage(sb, off)
        m.run_staged()
        ts = []
        for _ in range(a.reps):
            m.flush_l2()
            ts.append(m.run_staged())
        st = m.stats()
        best = min(t[0] for t in ts)
        print(c, f"pass {best:.3f} ms  kernel {min(t[1] for t in ts):.3f} ms  {a.segments / best / 1e3:.1f} M seg/s  cand/seg {st['seed_candidates'] / a.segments:.0f}", flush=True)
