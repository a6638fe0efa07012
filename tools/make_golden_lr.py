#!/usr/bin/env python3
"""Generate the extract_lr_bc fixtures under tests/golden/lr_bc/ by running the UNMODIFIED reference
(`/root/reference/scTagger.py extract_lr_bc`: read_fastqs, get_alns, get_ranges, filter_aln, output_matches, ...) with
ONE substitution: the `edlib` module, which is absent from this environment, is replaced by the pure-Python restatement
of `edlib.align(query, target, 'HW', 'locations')` below.  So these fixtures pin everything of
the stage against the reference's own code EXCEPT edlib's arithmetic, whose parity stays unpinned.

Only runs in the build container (needs /root/reference).   python tools/make_golden_lr.py
"""
import gzip
import importlib.util
import json
import multiprocessing
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sctagger_b200 import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "lr_bc")
REF = "/root/reference/scTagger.py"


def edlib_align(query, target, mode="HW", task="locations", k=-1):
    """Restated edlib semantics for the one call shape the reference uses (scTagger.py:180-181)."""
    assert mode == "HW" and task == "locations" and k == -1
    m, n = len(query), len(target)
    if n == 0:
        return {"editDistance": m, "locations": [(None, -1)]}
    col = list(range(m + 1))
    best, ends = None, []
    for j in range(1, n + 1):
        diag, col[0] = col[0], 0
        c = target[j - 1]
        for i in range(1, m + 1):
            v = min(diag + (query[i - 1] != c), col[i - 1] + 1, col[i] + 1)
            diag, col[i] = col[i], v
        if best is None or col[m] < best:
            best, ends = col[m], []
        if col[m] == best:
            ends.append(j - 1)
    locs = []
    rq = query[::-1]
    for e in ends:
        # reversed query against target[e], target[e-1], ...: prefix-anchored, take the LAST column with the best score
        col = list(range(m + 1))
        last = None
        for p in range(1, min(e + 1, m + best) + 1):
            diag, col[0] = col[0], p
            c = target[e - (p - 1)]
            for i in range(1, m + 1):
                v = min(diag + (rq[i - 1] != c), col[i - 1] + 1, col[i] + 1)
                diag, col[i] = col[i], v
            if col[m] == best:
                last = p - 1
        locs.append((e - last, e))
    return {"editDistance": best, "locations": locs}


def load_reference():
    for name in ("matplotlib", "matplotlib.ticker", "matplotlib.pyplot", "pysam", "ahocorasick"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["matplotlib"].ticker = sys.modules["matplotlib.ticker"]
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    ed = types.ModuleType("edlib")
    ed.align = edlib_align
    sys.modules["edlib"] = ed
    spec = importlib.util.spec_from_file_location("scTagger", REF)
    mod = importlib.util.module_from_spec(spec)
    sys.modules["scTagger"] = mod
    spec.loader.exec_module(mod)
    return mod


def make_inputs():
    os.makedirs(OUT, exist_ok=True)
    rng = np.random.default_rng(77)
    rb, off = synth.make_reads(260, 5001, median_len=220, min_len=60, max_len=500)
    rb = rb.copy()
    rb[rng.integers(0, rb.size, size=40)] = ord("N")
    rb[rng.integers(0, rb.size, size=40)] = ord("a")
    raw = rb.tobytes()
    reads = [raw[off[i]:off[i + 1]] for i in range(260)]
    reads[5] = b""                       # empty sequence line
    reads[6] = b"ACGT"                   # shorter than the adapter
    reads[7] = synth.ADAPTER             # the adapter alone
    reads[8] = b"N" * 40
    reads[9] = synth.ADAPTER + b"ACGTACGTAC" + synth.ADAPTER + b"TTTTTTTTTTTTTTTTTTTTTTTTTTTTTT"   # two exact hits

    def fastq(path, rs, first):
        opener = gzip.open if path.endswith(".gz") else open
        with opener(path, "wb") as f:
            for i, s in enumerate(rs):
                f.write(b"@r%d extra words here\n%s\n+\n%s\n" % (first + i, s, b"I" * len(s)))

    fastq(os.path.join(OUT, "readsA.fastq"), reads[:160], 0)
    fastq(os.path.join(OUT, "readsB.fastq.gz"), reads[160:], 160)


CASES = [
    {"name": "detect", "reads": ["readsA.fastq", "readsB.fastq.gz"], "args": ["-t", "2"]},
    {"name": "detect_bp25", "reads": ["readsA.fastq"], "args": ["-t", "1", "--num-bp-after", "25"]},
    {"name": "ranges", "reads": ["readsA.fastq", "readsB.fastq.gz"], "args": ["-t", "2", "-g", "f25:90", "r25:90"]},
    {"name": "ranges_two", "reads": ["readsB.fastq.gz"], "args": ["-t", "1", "-g", "f1:40", "f50:95", "r1:60"]},
    {"name": "adapter10", "reads": ["readsA.fastq"], "args": ["-t", "2", "-sa", "CGCTCTTCCG"]},
]


def main():
    multiprocessing.set_start_method("fork")
    make_inputs()
    mod = load_reference()
    for c in CASES:
        out = os.path.join(OUT, f"ref_{c['name']}.tsv.gz")
        sys.argv = [REF, "extract_lr_bc", "-r"] + [os.path.join(OUT, r) for r in c["reads"]] + c["args"] + ["-o", out]
        mod.main()
        c["out"] = os.path.basename(out)
        with gzip.open(out, "rb") as f:
            rows = f.read().splitlines()
        print(c["name"], len(rows), "rows,", sum(1 for r in rows if r.split(b"\t")[1] != b"-1"), "kept", file=sys.stderr)
    with open(os.path.join(OUT, "manifest.json"), "w") as f:
        json.dump({"cases": CASES, "note": "reference outputs with edlib replaced by the restatement in tools/make_golden_lr.py"}, f, indent=1)


if __name__ == "__main__":
    main()
