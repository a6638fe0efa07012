"""End-to-end wall clock of the `match_trie` CLI, -lr file.tsv.gz -> -o out.tsv.gz (SURVEY.md §8(d) figure (ii)), with the
phases timed in-process (read -sr, read -lr, GPU pass, format + gzip + write) and the whole command timed as a subprocess.

    python tools/cli_e2e.py [n_segments] [max_error] [n_gpus]      (run on the GPU box; files go to a temp dir)
"""
import gzip, hashlib, json, os, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from sctagger_b200 import _lib, synth, match_trie as mt

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
mr = int(sys.argv[2]) if len(sys.argv) > 2 else 2
gpus = int(sys.argv[3]) if len(sys.argv) > 3 else 1
threads = 16
out = {"segments": n, "barcodes": 25_000, "max_error": mr, "gpus": gpus, "host_cores": os.cpu_count()}
with tempfile.TemporaryDirectory() as d:
    bc = synth.make_barcodes(25_000, synth.SEEDS["cfg2"])
    sb, off = synth.make_segments(bc, n, 2002)
    lr, sr, o = os.path.join(d, "lr.tsv.gz"), os.path.join(d, "sr.tsv.gz"), os.path.join(d, "out.tsv.gz")
    synth.write_barcodes(sr, bc, 1)
    synth.write_segments(lr, sb, off)
    out["lr_gz_MB"] = os.path.getsize(lr) / 1e6
    del sb, off
    t = [time.perf_counter()]
    barcodes = mt.read_sr_barcodes(sr); t.append(time.perf_counter())
    reads = mt.read_long_reads(lr, os.cpu_count()); t.append(time.perf_counter())
    res = mt.run_get_matches_prefix(barcodes, reads, mr, threads, 16, 16.0, devices=gpus); t.append(time.perf_counter())
    res = mt.run_get_matches_prefix(barcodes, reads, mr, threads, 16, 16.0, devices=gpus); t.append(time.perf_counter())
    data = mt.format_reads_rows(reads, res, barcodes, gzip_level=mt.GZIP_LEVEL); t.append(time.perf_counter())
    with open(o, "wb") as f:
        f.write(data)
    t.append(time.perf_counter())
    out["phases_s"] = {"read_sr": t[1] - t[0], "read_lr": t[2] - t[1], "match_first_call_incl_cuda_init": t[3] - t[2],
                       "match_second_call": t[4] - t[3], "format_gzip": t[5] - t[4], "write": t[6] - t[5]}
    out["rows"] = int((res.min_ed >= 0).sum())
    out["out_gz_MB"] = len(data) / 1e6
    ref_digest = hashlib.sha256(gzip.decompress(data)).hexdigest()
    os.unlink(o)
    dev = ["--gpus", str(gpus)] if gpus > 1 else []
    t0 = time.perf_counter()
    subprocess.run([sys.executable, "-m", "sctagger_b200", "match_trie", "-lr", lr, "-sr", sr, "-mr", str(mr), "-t", str(threads),
                    "-o", o] + dev, check=True, cwd=ROOT, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    out["cli_wall_s"] = time.perf_counter() - t0
    with gzip.open(o, "rb") as f:
        out["cli_output_identical"] = hashlib.sha256(f.read()).hexdigest() == ref_digest
    out["cli_segments_per_s"] = n / out["cli_wall_s"]
print(json.dumps(out))
