"""BASELINE config 5 — extract_lr_bc adapter alignment — measured like bench.py measures match_trie (run on the GPU box):
    python tools/bench_align.py [--reads N] [--steps K] [--warmup W] [--cpu-reads M]
One JSON line: reads/s with the batch resident in HBM (CUDA events around both passes), e2e through Aligner.align
(host numpy in/out, H2D/D2H inside), algorithmic roofline (14 lane-ops x 2 adapters x sum(len)) against
the live-measured LOP3 peak, HBM GB/s as the minor counter, and the CPU bit-vector port on the host cores."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sctagger_b200 import _lib, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--reads", type=int, default=200_000)
ap.add_argument("--steps", type=int, default=10)
ap.add_argument("--warmup", type=int, default=3)
ap.add_argument("--cpu-reads", type=int, default=200_000)
ap.add_argument("--no-cpu", action="store_true")
a = ap.parse_args()

BLOCK = 200_000


def _block(b):
    return synth.make_reads(min(BLOCK, a.reads - b * BLOCK), synth.SEEDS["cfg5"] + b)


t0 = time.perf_counter()
n_blocks = (a.reads + BLOCK - 1) // BLOCK
if n_blocks == 1:
    parts = [_block(0)]
else:                                            # the 10 M-read configuration: 50 blocks of 200 k, seeds 1005 + b
    import multiprocessing as mp
    with mp.get_context("fork").Pool(min(os.cpu_count() or 1, n_blocks)) as pool:
        parts = pool.map(_block, range(n_blocks))
total_len = int(sum(int(o[-1]) for _, o in parts))
prb, poff, out = _lib.Pinned(total_len, np.uint8), _lib.Pinned(a.reads + 1, np.int64), _lib.AlignBuffers(a.reads)
rb, off = prb.array, poff.array                 # generated straight into page-locked memory (the reader's buffers in the CLI)
at_b = at_r = 0
off[0] = 0
for pb, po in parts:
    rb[at_b:at_b + pb.size] = pb
    off[at_r + 1:at_r + po.size] = po[1:] + at_b
    at_b += pb.size
    at_r += po.size - 1
del parts
gen_s = time.perf_counter() - t0
peak = _lib.measure_int32_peak(0, 0, 5)
with _lib.Aligner(synth.ADAPTER, threads=16, mem_gb=16.0) as al:
    al.stage(rb, off)
    for _ in range(a.warmup):
        al.run_staged()
    ms = k1 = 0.0
    for _ in range(a.steps):
        x, y = al.run_staged()
        ms += x
        k1 += y
    ms /= a.steps
    k1 /= a.steps
    for _ in range(2):
        res = al.align(rb, off, out=out, copy_locs=False)
    t0 = time.perf_counter()
    for _ in range(a.steps):
        res = al.align(rb, off, out=out, copy_locs=False)
    e2e_s = (time.perf_counter() - t0) / a.steps
    st = al.stats()
    pg = None
    if a.reads <= 1_000_000:                     # the same call on ordinary numpy arrays (staged by the library's host threads)
        rb2, off2 = np.array(rb), np.array(off)
        al.align(rb2, off2)
        t0 = time.perf_counter()
        for _ in range(a.steps):
            res_pg = al.align(rb2, off2)
        pg = (time.perf_counter() - t0) / a.steps
        assert all(np.array_equal(x, y) for x, y in zip(res, res_pg))
    res = tuple(np.array(x) for x in res)
work = 14.0 * 2.0 * total_len
line = {
    "metric": "extract_lr_bc_reads_per_sec_22nt_adapter", "workload": f"{a.reads:,} synthetic long reads (log-normal, median 900, "
    f"200..5000 nt, blocks of 200 k seeded 1005 + b) x adapter CTACACGACGCTCTTCCGATCT and its reverse complement (BASELINE configs[4] shape; the full "
    f"config is 10 M reads)", "value": a.reads / (ms * 1e-3), "unit": "reads/s", "ms_per_step": ms, "pass1_ms": k1,
    "bases_per_s": total_len / (ms * 1e-3),
    "e2e": {"value": a.reads / e2e_s, "unit": "reads/s", "ms_per_step": 1e3 * e2e_s, "h2d_bytes_per_step": st["h2d_bytes"],
            "d2h_bytes_per_step": st["d2h_bytes"], "host_GBps": st["h2d_bytes"] / e2e_s / 1e9, "chunks": st["chunks"],
            "api": "Aligner.align -> sct_align, page-locked host arrays in and out (no staging pass)",
            "pageable": None if pg is None else {"value": a.reads / pg, "ms_per_step": 1e3 * pg}},
    "generate_s": gen_s,
    "roofline": {"bound": "int32_alu", "achieved": work / (ms * 1e-3) / 1e12, "peak": peak / 1e12, "unit": "T lane-op/s",
                 "frac": work / (ms * 1e-3) / peak, "algorithmic_ops_per_launch": work,
                 "hbm_minor": {"GBps": (2 * total_len) / (ms * 1e-3) / 1e9, "note": "read bytes are streamed twice (two passes)"}},
    "check": {"aligned": int((res[1] >= 0).sum()), "plus": int((res[0] == ord('+')).sum()), "locs": int(res[3].size)},
}
if not a.no_cpu:
    from oracle import oracle
    n = min(a.cpu_reads, a.reads)
    cores = os.cpu_count() or 1
    t0 = time.perf_counter()
    want = oracle.align(synth.ADAPTER, rb[:off[n]], off[:n + 1], workers=cores, algo="bv")
    dt = time.perf_counter() - t0
    line["cpu_baseline"] = {"value": n / dt, "unit": "reads/s", "cores": cores, "kind": "port",
                            "sample": f"{n:,} reads, {dt:.2f} s, oracle/align_oracle.c (Myers bit-vector, what edlib runs) on {cores} threads"}
    line["check"]["equals_cpu"] = bool(all(np.array_equal(x[:len(y)] if i < 2 else x, y) for i, (x, y) in enumerate(
        zip((res[0][:n], res[1][:n]), (want[0], want[1])))))
print(json.dumps(line), flush=True)
