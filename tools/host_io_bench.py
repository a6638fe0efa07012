"""Host-side cost of the match_trie file formats (native sct_io_* vs the Python restatement): read -lr, format + gzip -o.
No GPU: the match result is synthetic (one tie per row).   python tools/host_io_bench.py [n_segments] [threads]"""
import gzip, os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from sctagger_b200 import synth, match_trie as mt
import host_model as hm

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
threads = int(sys.argv[2]) if len(sys.argv) > 2 else (os.cpu_count() or 1)
bc = synth.make_barcodes(25_000, 1002)
sb, off = synth.make_segments(bc, n, 2002)
with tempfile.TemporaryDirectory() as d:
    p = os.path.join(d, "lr.tsv.gz")
    synth.write_segments(p, sb, off)
    print(f"{n:,} segments, -lr {os.path.getsize(p) / 1e6:.1f} MB gz, {threads} threads")
    t0 = time.perf_counter(); reads = mt.read_long_reads(p, threads); t1 = time.perf_counter()
    print(f"native read_long_reads      {t1 - t0:8.3f} s")
    rng = np.random.default_rng(1)
    min_ed = np.where(rng.random(n) < 0.89, rng.integers(0, 3, n), -1).astype(np.int32)
    cnt = (min_ed >= 0).astype(np.int32)
    to = np.zeros(n + 1, np.int64); to[1:] = np.cumsum(cnt)
    res = mt.MatchResult(min_ed, cnt, to, rng.integers(0, 50_000, int(to[-1])).astype(np.int32))
    bcs = [bytes(b).decode() for b in bc]
    for level in (0, 1, 2, 3, 4, 5, 6):
        t0 = time.perf_counter(); out = mt.format_reads_rows(reads, res, bcs, gzip_level=level, threads=threads); t1 = time.perf_counter()
        print(f"native format level {level}       {t1 - t0:8.3f} s   {len(out) / 1e6:8.1f} MB")
    if n <= 200_000:
        t0 = time.perf_counter(); names, segs = hm.read_long_reads(p); t1 = time.perf_counter()
        print(f"python read_long_reads      {t1 - t0:8.3f} s")
        t0 = time.perf_counter(); plain = hm.format_matches(names, segs, res.min_ed, res.n_ties, res.tie_offsets, res.tie_ids, [bytes(b) for b in bc]); t1 = time.perf_counter()
        print(f"python format               {t1 - t0:8.3f} s")
        t0 = time.perf_counter(); z = gzip.compress(plain, 9); t1 = time.perf_counter()
        print(f"python gzip level 9         {t1 - t0:8.3f} s   {len(z) / 1e6:8.1f} MB")
        assert plain == mt.format_reads_rows(reads, res, bcs)
