#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY — runs the UNMODIFIED reference `scTagger.py` from /root/reference.

The reference imports matplotlib/edlib/pysam/ahocorasick at module top level
(reference scTagger.py:12-17); none of them is touched by `match_trie`
(scTagger.py:758-790), so empty stub modules are registered before the file is executed.
Workers are forked (`multiprocessing.Pool`, scTagger.py:703) and read module globals, so the
module is registered as `sys.modules['scTagger']` and the fork start method is forced.

Usage:  python oracle/ref_harness.py match_trie -lr L -sr S -mr 2 -t 16 -o out.tsv.gz
This file only works where /root/reference exists (the build container, never the GPU box);
it is used by tools/make_golden.py to create the fixtures committed under tests/golden/.
"""
import importlib.util
import multiprocessing
import os
import sys
import types

REF = os.environ.get("SCTAGGER_REFERENCE", "/root/reference/scTagger.py")


def load_reference():
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
