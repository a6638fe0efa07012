#!/usr/bin/env python3
"""Generate the golden fixtures under tests/golden/ by running the UNMODIFIED reference
(`/root/reference/scTagger.py match_trie`, via oracle/ref_harness.py) on seeded inputs.

Only runs in the build container (needs /root/reference).  Inputs AND reference outputs are
committed, together with this script, so that the GPU box (which has no /root/reference) can
check both the oracle and the CUDA path against the real reference's bytes.

    python tools/make_golden.py            # regenerate everything (~1-2 min)
"""
import gzip
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sctagger_b200 import synth  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
HARNESS = os.path.join(ROOT, "oracle", "ref_harness.py")


def rc(s):
    m = {"A": "T", "C": "G", "G": "C", "T": "A"}
    return "".join(m.get(c, c) for c in reversed(s))


def rand_dna(rng, n):
    return "".join("ACGT"[i] for i in rng.integers(0, 4, size=n))


def mutate(rng, s, e):
    s = list(s)
    for _ in range(e):
        k = rng.integers(0, 3)
        if k == 0 and s:
            s[rng.integers(0, len(s))] = "ACGT"[rng.integers(0, 4)]
        elif k == 1:
            s.insert(rng.integers(0, len(s) + 1), "ACGT"[rng.integers(0, 4)])
        elif s:
            del s[rng.integers(0, len(s))]
    return "".join(s)


def edge_case_inputs(seed, bl=16, n_random_bc=40, n_rows=360):
    """Barcodes and segments covering SURVEY.md Appendix A.5's edge list."""
    rng = np.random.default_rng(seed)
    unit = "ACGT" * 8
    bcs = [rand_dna(rng, bl) for _ in range(n_random_bc)]
    pal = unit[:bl] if bl % 2 == 0 else None  # ACGTACGT... is its own reverse complement for even bl%4==0
    if pal and rc(pal) == pal:
        bcs.append(pal)
    bcs.append(bcs[3])                      # duplicated barcode line
    nb = list(rand_dna(rng, bl)); nb[5 % bl] = "N"; bcs.append("".join(nb))   # barcode containing N
    bcs.append("A" * bl)                    # poly-A (N maps to A -> many ties)
    bcs.append("A" * (bl - 1) + "C")
    bcs.append("C" + "A" * (bl - 1))
    near = list(bcs[0]); near[7 % bl] = "A" if near[7 % bl] != "A" else "C"; bcs.append("".join(near))  # 1 sub from bc0
    segs = []
    segs += ["", "A", rand_dna(rng, bl - 3), rand_dna(rng, bl - 2), rand_dna(rng, bl - 1)]
    segs += [bcs[0], rc(bcs[1]), bcs[2][: bl - 2], bcs[2][1:], bcs[2][2:], bcs[4][:-1], bcs[4][:-2]]
    segs += [bcs[5] + rand_dna(rng, 6), rand_dna(rng, 6) + bcs[6], rand_dna(rng, 3) + rc(bcs[7]) + rand_dna(rng, 3)]
    if pal and rc(pal) == pal:
        segs += ["GG" + pal + "TT", pal]
    segs += ["TT" + bcs[3] + "GGGG", "A" * 22, "N" * 22, "A" * 10 + "N" * 4 + "A" * 10, "a" * 22]
    # N / lower-case inside, and in the first k bytes of a candidate start
    for b in (bcs[8], bcs[9], bcs[10]):
        s = "GCA" + b + "TGC"
        for pos in (0, 1, 2, 3, 4, 5, 10, len(s) - 4, len(s) - 3, len(s) - 2, len(s) - 1):
            t = list(s); t[pos] = "N"; segs.append("".join(t))
            t = list(s); t[pos] = t[pos].lower(); segs.append("".join(t))
        t = list(s); t[3] = "N"; t[4] = "N"; segs.append("".join(t))
        t = list(s); t[2] = "n"; t[3] = "n"; t[4] = "n"; segs.append("".join(t))
        segs.append(s.lower())
    # long segments (multi-window path), some dirty
    for L in (31, 32, 33, 40, 47, 64, 70, 100, 150, 400):
        b = bcs[int(rng.integers(0, n_random_bc))]
        core = mutate(rng, b if rng.random() < 0.5 else rc(b), int(rng.integers(0, 4)))
        p = int(rng.integers(0, max(1, L - len(core))))
        s = rand_dna(rng, p) + core + rand_dna(rng, max(0, L - p - len(core)))
        segs.append(s)
        t = list(s); t[int(rng.integers(0, len(t)))] = "N"; t[int(rng.integers(0, len(t)))] = "N"; segs.append("".join(t))
        # two different barcodes far apart in one long segment
        s2 = rand_dna(rng, 5) + bcs[11] + rand_dna(rng, max(0, L - 2 * bl - 10)) + mutate(rng, bcs[12], 1) + rand_dna(rng, 5)
        segs.append(s2)
    while len(segs) < n_rows:
        b = bcs[int(rng.integers(0, len(bcs)))]
        if rng.random() < 0.15:
            core = rand_dna(rng, bl)
        else:
            core = mutate(rng, b if rng.random() < 0.5 else rc(b), int(rng.integers(0, 5)))
        s = rand_dna(rng, int(rng.integers(0, 5))) + core + rand_dna(rng, int(rng.integers(0, 8)))
        if rng.random() < 0.15:
            t = list(s)
            if t:
                i = int(rng.integers(0, len(t)))
                t[i] = "N" if rng.random() < 0.6 else t[i].lower()
            s = "".join(t)
        segs.append(s)
    return bcs, segs


def write_inputs(dirpath, bcs, segs, gz_lr=True):
    os.makedirs(dirpath, exist_ok=True)
    with open(os.path.join(dirpath, "sr.tsv"), "w") as f:
        for i, b in enumerate(bcs):
            f.write(f"{b}\t{1000 - i}\n")
    lr = os.path.join(dirpath, "lr.tsv.gz" if gz_lr else "lr.tsv")
    opener = (lambda p: gzip.open(p, "wt", compresslevel=9)) if gz_lr else (lambda p: open(p, "w"))
    with opener(lr) as f:
        for i, s in enumerate(segs):
            d = i % 7
            f.write(f"r{i}/x\t{d}\t{i % 50}\t{s}\n")
    return os.path.join(dirpath, "sr.tsv"), lr


def run_ref(lr, sr, mr, t, bl, out):
    cmd = [sys.executable, HARNESS, "match_trie", "-lr", lr, "-sr", sr, "-mr", str(mr), "-t", str(t),
           "-bl", str(bl), "-o", out]
    subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    # store decompressed bytes re-compressed deterministically (gzip mtime makes raw bytes unstable)
    if out.endswith(".gz"):
        data = gzip.open(out, "rb").read()
        with open(out, "wb") as raw, gzip.GzipFile(fileobj=raw, mode="wb", compresslevel=9, mtime=0, filename="") as g:
            g.write(data)


def ties_inputs(seed=21, bl=16):
    """A barcode list made of the 1- and 2-substitution neighbourhood of one 16-mer: segments holding
    the centre tie against dozens (e=1) or ~1000 (e=2) barcodes -> exercises tie-slab overflow."""
    rng = np.random.default_rng(seed)
    x = rand_dna(rng, bl)
    bcs = []
    for i in range(bl):
        for c in "ACGT":
            if c != x[i]:
                bcs.append(x[:i] + c + x[i + 1:])
    one = len(bcs)
    for i in range(bl):
        for j in range(i + 1, bl):
            for c in "ACGT":
                for e in "ACGT":
                    if c != x[i] and e != x[j]:
                        bcs.append(x[:i] + c + x[i + 1:j] + e + x[j + 1:])
    order = rng.permutation(len(bcs))
    bcs = [bcs[i] for i in order]
    y = rand_dna(rng, bl)
    bcs.append(y)
    segs = ["GCA" + x + "TGC", "TT" + rc(x) + "GGGG", x, "GCA" + bcs[0] + "TGC", "GCA" + bcs[one + 5] + "TGC",
            "GCA" + x[:7] + x[8:] + "TGCA", "GCA" + x[:7] + "N" + x[8:] + "TGC", "GCA" + y + "TGC",
            rand_dna(rng, 22), "ACG" + mutate(rng, x, 3) + "TTA", "ACG" + mutate(rng, rc(x), 2) + "TTA",
            rand_dna(rng, 30) + x + rand_dna(rng, 30)]
    return bcs, segs


def main():
    manifest = []
    only = sys.argv[sys.argv.index("--only") + 1] if "--only" in sys.argv else None
    if only == "ties":
        d = os.path.join(GOLD, "ties16")
        bcs, segs = ties_inputs()
        sr, lr = write_inputs(d, bcs, segs, gz_lr=False)
        old = json.load(open(os.path.join(GOLD, "manifest.json")))
        cases = [c for c in old["cases"] if c["dir"] != "ties16"]
        for mr, t in ((1, 16), (2, 16), (2, 1), (3, 4)):
            out = os.path.join(d, f"ref_mr{mr}_t{t}.tsv.gz")
            run_ref(lr, sr, mr, t, 16, out)
            cases.append({"dir": "ties16", "lr": "lr.tsv", "sr": "sr.tsv", "mr": mr, "t": t, "bl": 16,
                          "out": os.path.basename(out)})
        old["cases"] = cases
        json.dump(old, open(os.path.join(GOLD, "manifest.json"), "w"), indent=1)
        print("wrote ties16")
        return
    # 1. edge-case fixture: bl=16, every mr in 0..3, every k-class of -t
    d = os.path.join(GOLD, "edge16")
    bcs, segs = edge_case_inputs(seed=11)
    sr, lr = write_inputs(d, bcs, segs)
    for mr in (0, 1, 2, 3):
        for t in (1, 4, 5, 16, 17):
            out = os.path.join(d, f"ref_mr{mr}_t{t}.tsv.gz")
            run_ref(lr, sr, mr, t, 16, out)
            manifest.append({"dir": "edge16", "lr": "lr.tsv.gz", "sr": "sr.tsv", "mr": mr, "t": t, "bl": 16,
                             "out": os.path.basename(out)})
    # 2. short barcodes
    for bl, mrs, seed in ((12, (1, 2), 12), (8, (0, 1, 2), 13), (10, (3,), 14)):
        d = os.path.join(GOLD, f"edge{bl}")
        bcs, segs = edge_case_inputs(seed=seed, bl=bl, n_random_bc=30, n_rows=260)
        sr, lr = write_inputs(d, bcs, segs, gz_lr=False)
        for mr in mrs:
            for t in (1, 16):
                out = os.path.join(d, f"ref_mr{mr}_t{t}.tsv")
                run_ref(lr, sr, mr, t, bl, out)
                manifest.append({"dir": f"edge{bl}", "lr": "lr.tsv", "sr": "sr.tsv", "mr": mr, "t": t, "bl": bl,
                                 "out": os.path.basename(out)})
    # 3. BASELINE config #1: 10k synthetic segments x 1k barcodes (+1 % edge rows), -mr 2 -t 16
    d = os.path.join(GOLD, "cfg1")
    os.makedirs(d, exist_ok=True)
    seed = synth.SEEDS["cfg1"]
    bc = synth.make_barcodes(1000, seed)
    sb, off = synth.make_segments(bc, 10000, seed)
    bcs = [r.tobytes().decode() for r in bc]
    raw = sb.tobytes()
    segs = [raw[off[i]:off[i + 1]].decode() for i in range(10000)]
    _, extra = edge_case_inputs(seed=15, n_random_bc=20, n_rows=100)
    rng = np.random.default_rng(16)
    for s in extra[:100]:
        # re-target some edge rows at real cfg1 barcodes so they produce rows
        segs[int(rng.integers(0, len(segs)))] = s
    for j in range(40):
        b = bcs[int(rng.integers(0, 1000))]
        t = list("GCA" + b + "TGC"); t[int(rng.integers(0, len(t)))] = "N"
        segs[int(rng.integers(0, len(segs)))] = "".join(t)
    with open(os.path.join(d, "sr.tsv"), "w") as f:
        cnt = synth.barcode_counts(1000, seed)
        for b, c in zip(bcs, cnt):
            f.write(f"{b}\t{c}\n")
    with gzip.open(os.path.join(d, "lr.tsv.gz"), "wt", compresslevel=9) as f:
        for i, s in enumerate(segs):
            f.write(f"read{i}\t0\t0\t{s}\n")
    for mr, t in ((2, 16), (3, 16), (2, 1)):
        out = os.path.join(d, f"ref_mr{mr}_t{t}.tsv.gz")
        run_ref(os.path.join(d, "lr.tsv.gz"), os.path.join(d, "sr.tsv"), mr, t, 16, out)
        manifest.append({"dir": "cfg1", "lr": "lr.tsv.gz", "sr": "sr.tsv", "mr": mr, "t": t, "bl": 16,
                         "out": os.path.basename(out)})
    with open(os.path.join(GOLD, "manifest.json"), "w") as f:
        json.dump({"generator": "tools/make_golden.py", "reference": "vpc-ccg/scTagger scTagger.py match_trie (unmodified)",
                   "cases": manifest}, f, indent=1)
    print(f"wrote {len(manifest)} golden cases")


if __name__ == "__main__":
    main()
