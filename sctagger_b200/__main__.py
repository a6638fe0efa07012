"""`python -m sctagger_b200 match_trie ...` == `scTagger.py match_trie ...` (reference scTagger.py:864-865)."""
import sys

from .cli import main

sys.exit(main())
