#!/bin/bash
# TEST INFRASTRUCTURE ONLY — "installs" the UNMODIFIED reference for timing on the GPU box.
# The reference is a single Python script (no build system, nothing to compile): the analogue of oracle/_ref/*.so for a
# compiled reference is a byte-identical copy of /root/reference/scTagger.py under oracle/_ref/ (git-ignored, so it never
# enters history; NOT gpurun-ignored, so it travels to the GPU box like the built .so files).  oracle/ref_harness.py
# executes it unmodified (it only stubs the four third-party imports match_trie never touches); bench.py --impl reference
# and tools/time_reference.py time it on the box's host cores.
# Only runs where /root/reference exists (the build container); a no-op elsewhere.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
SRC="${SCTAGGER_REFERENCE_DIR:-/root/reference}/scTagger.py"
if [ ! -f "$SRC" ]; then
    echo "make_ref: $SRC not present (GPU box?) - keeping whatever is in $HERE/_ref" >&2
    exit 0
fi
mkdir -p "$HERE/_ref"
cp -f "$SRC" "$HERE/_ref/scTagger.py"
sha256sum "$HERE/_ref/scTagger.py" | cut -d' ' -f1 > "$HERE/_ref/scTagger.py.sha256"
echo "make_ref: $HERE/_ref/scTagger.py ($(cat "$HERE/_ref/scTagger.py.sha256"))"
