#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY — runs the UNMODIFIED reference `scTagger.py` from /root/reference.

The reference imports matplotlib/edlib/pysam/ahocorasick at module top level
(reference scTagger.py:12-17); none of them is touched by `match_trie`
(scTagger.py:758-790), so empty stub modules are registered before the file is executed.
Workers are forked (`multiprocessing.Pool`, scTagger.py:703) and read module globals, so the
module is registered as `sys.modules['scTagger']` and the fork start method is forced.

Usage:  python oracle/ref_harness.py match_trie -lr L -sr S -mr 2 -t 16 -o out.tsv.gz
It runs where /root/reference exists (the build container) or where oracle/make_ref.sh left its byte-identical copy under
oracle/_ref/ (the GPU box: timing the reference on the box's host cores, bench.py --impl reference and
tools/time_reference.py).  tools/make_golden.py uses it to create the fixtures committed under tests/golden/.
"""
import importlib.util
import multiprocessing
import os
import sys
import types

_HERE = os.path.dirname(os.path.abspath(__file__))


def _find_reference():
    """/root/reference in the build container; on the GPU box the byte-identical copy oracle/make_ref.sh put under
    oracle/_ref/ (untracked, travels with the snapshot)."""
    for p in (os.environ.get("SCTAGGER_REFERENCE"), "/root/reference/scTagger.py", os.path.join(_HERE, "_ref", "scTagger.py")):
        if p and os.path.exists(p):
            return p
    return None


REF = _find_reference()


def available() -> bool:
    return REF is not None


def load_reference():
    if REF is None:
        raise FileNotFoundError("the reference script is neither under /root/reference nor under oracle/_ref/ "
                                "(run oracle/make_ref.sh in the build container)")
    for name in ("matplotlib", "matplotlib.ticker", "matplotlib.pyplot", "edlib", "pysam", "ahocorasick"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["matplotlib"].ticker = sys.modules["matplotlib.ticker"]
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    spec = importlib.util.spec_from_file_location("scTagger", REF)
    mod = importlib.util.module_from_spec(spec)
    sys.modules["scTagger"] = mod
    spec.loader.exec_module(mod)
    return mod


def main(argv):
    try:
        multiprocessing.set_start_method("fork")
    except RuntimeError:
        pass
    mod = load_reference()
    sys.argv = [REF] + list(argv)
    mod.main()


if __name__ == "__main__":
    main(sys.argv[1:])
