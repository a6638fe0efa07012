#!/usr/bin/env python3
"""Time the UNMODIFIED reference `scTagger.py match_trie` on this machine's host cores:
`-t $(nproc)`, same synthetic files as the GPU arm, wall clock of the whole CLI.  Needs oracle/_ref/scTagger.py
(oracle/make_ref.sh) or /root/reference.

    python tools/time_reference.py cfg1                 # 10 k x 1 k, -mr 2          (~10 s)
    python tools/time_reference.py 100k [--check]       # 100 k x 25 k, -mr 2        (minutes); --check compares the digest
                                                        # with tests/golden/ref100k.json and with the CUDA path if a GPU is here
"""
import argparse, gzip, hashlib, json, os, subprocess, sys, tempfile, time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sctagger_b200 import synth  # noqa: E402

CASES = {"cfg1": (10_000, 1_000, 2, "cfg1", 2001), "100k": (100_000, 25_000, 2, "cfg2", 2002)}


def run(case, threads=None, check=False):
    n, B, mr, bseed, sseed = CASES[case]
    threads = threads or os.cpu_count() or 1
    with tempfile.TemporaryDirectory() as d:
        bc = synth.make_barcodes(B, synth.SEEDS[bseed])
        sb, off = synth.make_segments(bc, n, sseed)
        lr, sr, out = (os.path.join(d, x) for x in ("lr.tsv.gz", "sr.tsv.gz", "out.tsv.gz"))
        synth.write_barcodes(sr, bc, 1)
        synth.write_segments(lr, sb, off)
        t0 = time.time()
        subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "ref_harness.py"), "match_trie", "-lr", lr, "-sr", sr,
                        "-mr", str(mr), "-t", str(threads), "-o", out], check=True, stdout=subprocess.DEVNULL)
        secs = time.time() - t0
        data = gzip.open(out, "rb").read()
    rec = {"what": "unmodified reference scTagger.py match_trie (oracle/ref_harness.py)", "case": case, "segments": n,
           "barcodes": B, "max_error": mr, "threads_flag": threads, "host_cores": os.cpu_count(), "seconds": round(secs, 2),
           "segments_per_s": round(n / secs, 1), "rows": data.count(b"\n"), "sha256": hashlib.sha256(data).hexdigest()}
    if check and case == "100k":
        gold = json.load(open(os.path.join(ROOT, "tests", "golden", "ref100k.json")))
        rec["equals_committed_golden_digest"] = gold["sha256"] == rec["sha256"]
    return rec


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("case", choices=sorted(CASES))
    ap.add_argument("-t", type=int, default=None)
    ap.add_argument("--check", action="store_true")
    a = ap.parse_args()
    print(json.dumps(run(a.case, a.t, a.check)))
