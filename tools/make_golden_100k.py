#!/usr/bin/env python3
"""Pin the oracle and the CUDA path against the UNMODIFIED reference at the headline table size: 100,000 synthetic segments
x 25,000 barcodes (BASELINE configs[1] shape, generator seeds 1002 / 2002), -mr 2.  The reference needs ~200 s on 8 cores
for this; its decompressed output is ~7 MB, so only its SHA-256, row count and a few sample rows are committed
(tests/golden/ref100k.json).  The tests regenerate the input from the seeds and compare digests.

Only runs in the build container (needs /root/reference).   python tools/make_golden_100k.py
"""
import gzip, hashlib, json, os, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sctagger_b200 import synth  # noqa: E402

N, B, MR, T = 100_000, 25_000, 2, 8
NAME = "ref100k.json"
if len(sys.argv) > 3:      # python tools/make_golden_100k.py <segments> <max_error> <file name under tests/golden>
    N, MR, NAME = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
with tempfile.TemporaryDirectory() as d:
    bc = synth.make_barcodes(B, synth.SEEDS["cfg2"])
    sb, off = synth.make_segments(bc, N, 2002)
    lr, sr, out = os.path.join(d, "lr.tsv.gz"), os.path.join(d, "sr.tsv.gz"), os.path.join(d, "out.tsv.gz")
    synth.write_barcodes(sr, bc, 1)
    synth.write_segments(lr, sb, off)
    t0 = time.time()
    subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "ref_harness.py"), "match_trie", "-lr", lr, "-sr", sr,
                    "-mr", str(MR), "-t", str(T), "-o", out], check=True, stdout=subprocess.DEVNULL)
    secs = time.time() - t0
    data = gzip.open(out, "rb").read()
rows = data.split(b"\n")
meta = {"segments": N, "barcodes": B, "max_error": MR, "threads": T, "barcode_seed": synth.SEEDS["cfg2"], "segment_seed": 2002,
        "rows": len(rows) - 1, "bytes": len(data), "sha256": hashlib.sha256(data).hexdigest(),
        "first_rows": [r.decode() for r in rows[:3]], "reference_seconds": round(secs, 1), "host_cores": os.cpu_count(),
        "generator": "tools/make_golden_100k.py"}
json.dump(meta, open(os.path.join(ROOT, "tests", "golden", NAME), "w"), indent=1)
print(json.dumps(meta))
