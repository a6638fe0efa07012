"""Shared helpers for the parity tests: golden-fixture loading and oracle access."""
import gzip
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def read_maybe_gz(path):
    with open(path, "rb") as f:
        head = f.read(2)
    if head == b"\x1f\x8b":
        with gzip.open(path, "rb") as f:
            return f.read()
    with open(path, "rb") as f:
        return f.read()


def golden_cases():
    with open(os.path.join(GOLD, "manifest.json")) as f:
        return json.load(f)["cases"]


def case_id(c):
    return f"{c['dir']}-mr{c['mr']}-t{c['t']}"


def load_case(c):
    """-> (barcodes list[str], LongReads, expected output bytes)"""
    from sctagger_b200 import match_trie as mt
    d = os.path.join(GOLD, c["dir"])
    import contextlib
    import io
    with contextlib.redirect_stdout(io.StringIO()):
        barcodes = mt.read_sr_barcodes(os.path.join(d, c["sr"]))
        reads = mt.read_long_reads(os.path.join(d, c["lr"]))
    expected = read_maybe_gz(os.path.join(d, c["out"]))
    return barcodes, reads, expected


def result_from_tuple(t):
    from sctagger_b200.match_trie import MatchResult
    return MatchResult(*t)


def random_case(rng, n_bc, n_seg, bl=16, dirty=0.0, max_len=40, lower=False):
    """Random barcodes + segments derived from them (0-4 edits), optionally with non-ACGT bytes."""
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)
    bc = acgt[rng.integers(0, 4, size=(n_bc, bl))]
    comp = {65: 84, 67: 71, 71: 67, 84: 65}
    segs = []
    for _ in range(n_seg):
        r = rng.random()
        if r < 0.05:
            s = list(acgt[rng.integers(0, 4, size=int(rng.integers(0, bl + 3)))])
        else:
            core = list(bc[int(rng.integers(0, n_bc))]) if r < 0.85 else list(acgt[rng.integers(0, 4, size=bl)])
            for _e in range(int(rng.integers(0, 5))):
                k = int(rng.integers(0, 3))
                if k == 0 and core:
                    core[int(rng.integers(0, len(core)))] = int(acgt[rng.integers(0, 4)])
                elif k == 1:
                    core.insert(int(rng.integers(0, len(core) + 1)), int(acgt[rng.integers(0, 4)]))
                elif core:
                    del core[int(rng.integers(0, len(core)))]
            if rng.random() < 0.5:
                core = [comp[int(x)] for x in reversed(core)]
            left = list(acgt[rng.integers(0, 4, size=int(rng.integers(0, 6)))])
            room = max(0, max_len - len(left) - len(core))
            right = list(acgt[rng.integers(0, 4, size=int(rng.integers(0, min(room, 12) + 1)))])
            s = left + core + right
        if dirty and s and rng.random() < dirty:
            for _d in range(int(rng.integers(1, 3))):
                i = int(rng.integers(0, len(s)))
                s[i] = ord("N") if (not lower or rng.random() < 0.6) else (int(s[i]) | 0x20)
        segs.append(bytes(int(x) for x in s))
    offsets = np.zeros(n_seg + 1, dtype=np.int64)
    np.cumsum([len(s) for s in segs], out=offsets[1:])
    seg_bytes = np.frombuffer(b"".join(segs), dtype=np.uint8).copy()
    return bc, seg_bytes, offsets


def long_case(rng, n_bc, n_seg, bl=16, min_len=20, max_len=200, dirty=0.0, plant=2):
    """Random barcodes + LONG random segments with up to `plant` mutated barcodes (either strand) planted anywhere,
    optionally with non-ACGT bytes: exercises the windowed seeded kernel (window seams, start restrictions)."""
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)
    bc = acgt[rng.integers(0, 4, size=(n_bc, bl))]
    comp = {65: 84, 67: 71, 71: 67, 84: 65}
    segs = []
    for _ in range(n_seg):
        L = int(rng.integers(min_len, max_len + 1))
        s = list(acgt[rng.integers(0, 4, size=L)])
        for _p in range(int(rng.integers(0, plant + 1))):
            core = [int(x) for x in bc[int(rng.integers(0, n_bc))]]
            for _e in range(int(rng.integers(0, 4))):
                k = int(rng.integers(0, 3))
                if k == 0:
                    core[int(rng.integers(0, len(core)))] = int(acgt[rng.integers(0, 4)])
                elif k == 1:
                    core.insert(int(rng.integers(0, len(core) + 1)), int(acgt[rng.integers(0, 4)]))
                else:
                    del core[int(rng.integers(0, len(core)))]
            if rng.random() < 0.5:
                core = [comp[x] for x in reversed(core)]
            if len(core) <= L:
                at = int(rng.integers(0, L - len(core) + 1))
                s[at:at + len(core)] = core
        if dirty and rng.random() < dirty:
            for _d in range(int(rng.integers(1, 4))):
                i = int(rng.integers(0, len(s)))
                s[i] = ord("N") if rng.random() < 0.7 else (int(s[i]) | 0x20)
        segs.append(bytes(int(x) for x in s))
    offsets = np.zeros(n_seg + 1, dtype=np.int64)
    np.cumsum([len(s) for s in segs], out=offsets[1:])
    seg_bytes = np.frombuffer(b"".join(segs), dtype=np.uint8).copy()
    return bc, seg_bytes, offsets
