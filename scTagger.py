#!/usr/bin/env python3
"""Drop-in entry script: the reference's command line, literally —

    ./scTagger.py match_trie -lr L.tsv.gz -sr S.tsv.gz -mr 2 -bl 16 -t 16 -m 16 -o out.tsv.gz
    ./scTagger.py extract_lr_bc ... | extract_sr_bc_from_lr ...

(reference scTagger.py:25-137 parse_args, 839-865 main).  Everything is in sctagger_b200.cli; this file only gives the
package the reference's script name.  The stages run on the GPU through libsctagger_b200.so; there is no CPU fallback."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from sctagger_b200.cli import main  # noqa: E402

if __name__ == "__main__":
    sys.exit(main())
