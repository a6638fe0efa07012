"""sctagger_b200 — B200-native replacement of scTagger's `match_trie` hot path.

Host side (this package) mirrors the reference's stage interface
(`scTagger.py match_trie`, reference scTagger.py:81-97,758-790) and calls a C-ABI
shared library (`libsctagger_b200.so`, see include/sctagger_b200.h) through ctypes.
There is no CPU fallback: every compute entry point raises if the CUDA library is missing.
"""
__version__ = "0.1.0"
