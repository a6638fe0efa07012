"""e2e of Aligner.align vs chunk size (run on the GPU box)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sctagger_b200 import _lib, synth
rb, off = synth.make_reads(200_000, 1005)
for cb in [int(x) << 20 for x in sys.argv[1:]] or [256 << 20]:
    with _lib.Aligner(synth.ADAPTER, threads=16, chunk_bytes=cb) as al:
        for _ in range(2):
            al.align(rb, off)
        t0 = time.perf_counter()
        for _ in range(5):
            al.align(rb, off)
        dt = (time.perf_counter() - t0) / 5
        st = al.stats()
        print(f"chunk {cb >> 20} MiB: {1e3 * dt:.2f} ms  {200_000 / dt / 1e6:.1f} M reads/s  {st['h2d_bytes'] / dt / 1e9:.1f} GB/s  chunks {st['chunks']} device {st['device_ns'] / 1e6:.2f} ms", flush=True)
