"""Sweep of the streaming parameters of sct_match on one GPU (run on the GPU box): e2e ms per call with page-locked
buffers for the whole 10 M input and for one 1.25 M shard (what a rank holds at 8 GPUs), against the resident pass."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from sctagger_b200 import _lib, synth

bc = synth.make_barcodes(25_000, 1002)
mr = int(sys.argv[1]) if len(sys.argv) > 1 else 2
for n in (10_000_000, 1_250_000):
    sb, off = synth.make_segments_range(bc, 0, n)
    psb, poff, out = _lib.Pinned.copy_of(sb), _lib.Pinned.copy_of(off), _lib.ResultBuffers(n)
    with _lib.Matcher(bc, mr, 16) as m:
        m.stage(sb, off)
        m.run_staged()
        res = min(m.run_staged()[0] for _ in range(3))
    for stream, ramp, chain in ((1 << 20, 1 << 15, 0), (1 << 20, 1 << 15, 1 << 14), (1 << 20, 1 << 15, 1 << 15), (1 << 20, 1 << 15, 1 << 16), (1 << 20, 1 << 15, 1 << 17), (1 << 20, 1 << 18, 1 << 16)):
        with _lib.Matcher(bc, mr, 16, stream_segments=stream, ramp_segments=ramp, tail_segments=chain) as m:
            for _ in range(2):
                m.match(psb.array, poff.array, out=out, copy_ties=False)
            ts = []
            for _ in range(6):
                t0 = time.perf_counter()
                m.match(psb.array, poff.array, out=out, copy_ties=False)
                ts.append(time.perf_counter() - t0)
            print(json.dumps({"n": n, "resident_ms": round(res, 3), "stream": stream, "ramp": ramp, "tail": chain, "chunks": m.stats()["chunks"],
                              "e2e_ms_min": round(1e3 * min(ts), 3), "e2e_ms_med": round(1e3 * sorted(ts)[3], 3)}), flush=True)
    for x in (psb, poff, out):
        x.close()
