#!/usr/bin/env python3
"""Generate the extract_sr_bc_from_lr fixtures under tests/golden/wl/ by running the UNMODIFIED reference
(`/root/reference/scTagger.py extract_sr_bc_from_lr`: whitelist reader, Counter loop, sort, threshold cut, writer) with
ONE substitution: the `ahocorasick` module, absent from this environment, is replaced by oracle/wl_oracle.Automaton
(dictionary lookup with pyahocorasick's documented add_word / iter semantics).  The fixtures therefore pin everything of
the stage against the reference's own code EXCEPT pyahocorasick itself, whose parity stays unpinned.

Only runs in the build container (needs /root/reference).   python tools/make_golden_wl.py
"""
import contextlib
import gzip
import io
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_harness, wl_oracle  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "wl")


def rand_seq(rng, n):
    return "".join("ACGT"[i] for i in rng.integers(0, 4, n))


def make_inputs(rng, n_words, n_rows, trailing_newline=True):
    words = [rand_seq(rng, 16) for _ in range(n_words)]
    words[5] = wl_oracle.rev_compl(words[2])          # an RC pair inside the whitelist: the RC call overwrites
    words[9] = "ACGTACGTACGTACGT"                     # its own reverse complement
    words[11] = words[3]                              # duplicated line: the later index wins
    words[13] = "ACGTNNACGTACGTAA"                    # only matches literally
    words[0] = rand_seq(rng, 16)                      # index 0: -0 == 0
    rows = []
    weights = rng.random(n_words) ** 6                # skewed barcode usage so the threshold cut has something to cut
    weights /= weights.sum()
    for i in range(n_rows):
        kind = rng.random()
        w = words[rng.choice(n_words, p=weights)]
        if rng.random() < 0.5:
            w = wl_oracle.rev_compl(w)
        if kind < 0.08:
            rows.append(f"r{i}\t-1\tNA\t")
            continue
        if kind < 0.60:
            seg = rand_seq(rng, int(rng.integers(0, 5))) + w + rand_seq(rng, int(rng.integers(0, 8)))
        elif kind < 0.70:                             # two occurrences -> skipped
            seg = w + rand_seq(rng, 2) + words[int(rng.integers(0, n_words))]
        elif kind < 0.75:                             # one error -> no exact occurrence
            k = int(rng.integers(0, 16))
            seg = "AA" + w[:k] + ("A" if w[k] != "A" else "C") + w[k + 1:] + "TT"
        elif kind < 0.80:
            seg = "AC" + words[13] + "GT" if rng.random() < 0.5 else "ac" + w.lower()
        elif kind < 0.83:
            seg = w[:15]
        else:
            seg = rand_seq(rng, int(rng.integers(0, 40)))
        rows.append(f"r{i}\t{int(rng.integers(0, 4))}\t{int(rng.integers(-60, 60))}\t{seg}")
    wl_text = "\n".join(words) + ("\n" if trailing_newline else "")
    return wl_text, "\n".join(rows) + "\n"


def run_reference(argv):
    mod = ref_harness.load_reference()
    sys.modules["ahocorasick"].Automaton = wl_oracle.Automaton
    mod.ahocorasick = sys.modules["ahocorasick"]
    sys.argv = [ref_harness.REF] + argv
    with contextlib.redirect_stdout(io.StringIO()), contextlib.redirect_stderr(io.StringIO()):
        mod.main()


def main():
    os.makedirs(OUT, exist_ok=True)
    rng = np.random.default_rng(77)
    cases = []
    for name, n_words, n_rows, trail, gz in (("a", 300, 4000, True, True), ("b", 60, 1500, False, False)):
        wl_text, lr_text = make_inputs(rng, n_words, n_rows, trail)
        wl = f"{name}_whitelist.txt" + (".gz" if gz else "")
        lr = f"{name}_lr_bc.tsv" + (".gz" if gz else "")
        for path, text in ((wl, wl_text), (lr, lr_text)):
            with (gzip.open if gz else open)(os.path.join(OUT, path), "wb") as f:
                f.write(text.encode())
        for tag, extra in (("default", []), ("step20", ["--step-size", "20", "--thresh", "0.05"]),
                           ("cap7", ["--max-barcode-cnt", "7", "--step-size", "3"]),
                           ("thresh1", ["--thresh", "1.0", "--step-size", "5"])):
            out = f"ref_{name}_{tag}.tsv.gz"
            run_reference(["extract_sr_bc_from_lr", "-i", os.path.join(OUT, lr), "-wl", os.path.join(OUT, wl),
                           "-o", os.path.join(OUT, out)] + extra)
            cases.append({"wl": wl, "lr": lr, "args": extra, "out": out})
    with open(os.path.join(OUT, "manifest.json"), "w") as f:
        json.dump({"cases": cases, "generator": "tools/make_golden_wl.py"}, f, indent=1)
    print(f"wrote {len(cases)} cases to {OUT}")


if __name__ == "__main__":
    main()
