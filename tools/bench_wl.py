"""extract_sr_bc_from_lr whitelist search on B200: device-resident kernel time against the HBM roofline, end-to-end
through Whitelist.count (host buffers in), and the CPU restatement on a sample.   python tools/bench_wl.py [W] [n] [steps]
Workload: W random 16-mers (default 6.8 M = the size of the 10x 3M-february-2018 whitelist), n segments of 38 bases
(= 2 + 16 + 20, what extract_lr_bc cuts with --num-bp-after 20), 80 % holding the barcode of one of 10,000 cells (or its
reverse complement), the rest random."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from sctagger_b200 import _lib

W = int(sys.argv[1]) if len(sys.argv) > 1 else 6_800_000
n = int(sys.argv[2]) if len(sys.argv) > 2 else 4_000_000
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 10
SEG = 38
rng = np.random.default_rng(1006)
lut = np.frombuffer(b"ACGT", np.uint8)
wm = rng.integers(0, 4, size=(W, 16), dtype=np.uint8)
wbytes, woff = lut[wm].reshape(-1), np.arange(W + 1, dtype=np.int64) * 16
seg = rng.integers(0, 4, size=(n, SEG), dtype=np.uint8)
cells = rng.choice(W, size=min(W, 10_000), replace=False)      # a real sample holds thousands of cells, not millions
pick = cells[rng.integers(0, cells.shape[0], n)]
has = rng.random(n) < 0.8
rc = rng.random(n) < 0.5
core = np.where(rc[:, None], 3 - wm[pick][:, ::-1], wm[pick])
seg[has, 2:18] = core[has]
sb, so = lut[seg].reshape(-1), np.arange(n + 1, dtype=np.int64) * SEG
out = {"whitelist_words": W, "segments": n, "segment_length": SEG}
t0 = time.perf_counter()
wl = _lib.Whitelist(wbytes, woff)
out["create_s"] = time.perf_counter() - t0
st = wl.stats()
out["table"] = {"keys": st["table_keys"], "slots": st["table_slots"], "MiB": st["table_slots"] * 16 / 2**20}
stage_n = min(n, 2_000_000)
wl.stage(sb[:stage_n * SEG], so[:stage_n + 1])
for _ in range(3):
    wl.run_staged()
ms = [wl.run_staged() for _ in range(steps)]
k_ms = float(np.mean(ms))
windows = stage_n * (SEG - 16 + 1)
alg = windows * 32.0 + stage_n * (SEG + 8.0) + stage_n * 0.8 * 12.0     # one 32 B sector per probe + text + offsets + count/first atomics
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs") if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else None
out["kernel"] = {"segments": stage_n, "ms": k_ms, "ms_min": float(min(ms)), "segments_per_s": stage_n / (k_ms * 1e-3),
                 "probes_per_s": windows / (k_ms * 1e-3), "algorithmic_bytes": alg,
                 "roofline": {"bound": "hbm", "achieved": alg / (k_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                              "frac": (alg / (k_ms * 1e-3) / 1e9 / peak) if peak else None}}
wl.reset()
wl.count(sb, so)
t0 = time.perf_counter()
for _ in range(3):
    wl.reset()
    wl.count(sb, so)
    idx, cnt = wl.result()
e2e = (time.perf_counter() - t0) / 3
st = wl.stats()
out["e2e_pageable"] = {"s": e2e, "segments_per_s": n / e2e, "h2d_bytes": st["h2d_bytes"], "unique_barcodes": int(idx.shape[0]),
                       "counted_segments": int(cnt.sum())}
# the same calls on page-locked host arrays (sct_host_alloc): copied from directly, no staging pass
psb, pso = _lib.Pinned.copy_of(sb), _lib.Pinned.copy_of(so)
wl.reset()
wl.count(psb.array, pso.array)
t0 = time.perf_counter()
for _ in range(3):
    wl.reset()
    wl.count(psb.array, pso.array)
    idx2, cnt2 = wl.result()
e2e = (time.perf_counter() - t0) / 3
st = wl.stats()
assert np.array_equal(idx, idx2) and np.array_equal(cnt, cnt2)
out["e2e"] = {"s": e2e, "segments_per_s": n / e2e, "h2d_bytes": st["h2d_bytes"], "d2h_bytes": st["d2h_bytes"],
              "unique_barcodes": int(idx2.shape[0]), "counted_segments": int(cnt2.sum()),
              "api": "Whitelist.reset + count + result on page-locked host arrays"}
psb.close(); pso.close()
wl.close()
# CPU restatement (dictionary lookup per window, one core) on a sample
sample = 200_000
table = {}
raw_w = wbytes.tobytes()
t0 = time.perf_counter()
for i in range(W if W <= 1_000_000 else 1_000_000):
    table[raw_w[16 * i:16 * i + 16]] = i
build = time.perf_counter() - t0
raw = sb[:sample * SEG].tobytes()
t0 = time.perf_counter()
hits = 0
for r in range(sample):
    s = raw[r * SEG:(r + 1) * SEG]
    for j in range(SEG - 15):
        if s[j:j + 16] in table:
            hits += 1
dt = time.perf_counter() - t0
out["cpu_python_dict"] = {"sample_segments": sample, "segments_per_s": sample / dt, "cores": 1,
                          "note": "pure-Python window lookup, forward words of the first 1 M barcodes only (a lower bound of the work)"}
print(json.dumps(out))
