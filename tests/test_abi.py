"""CPU suite: the C-ABI library loads, exports every symbol include/sctagger_b200.h declares, and — with no GPU —
fails loudly instead of falling back to a CPU path."""
import os
import re

import numpy as np
import pytest

import util
from sctagger_b200 import _lib


def _declared_symbols():
    with open(os.path.join(util.ROOT, "include", "sctagger_b200.h")) as f:
        text = f.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sct_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    L = _lib.lib()
    declared = _declared_symbols()
    assert declared, "no prototypes parsed from the header"
    for name in declared:
        assert hasattr(L, name), f"{name} declared in include/sctagger_b200.h but not exported"
    assert sorted(_lib.EXPORTS) == declared


def test_product_never_references_the_oracle():
    """Nothing under sctagger_b200/ may import, load or link the oracle (it is test infrastructure)."""
    pkg = os.path.join(util.ROOT, "sctagger_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".sh", ".cpp")):
                with open(os.path.join(dirpath, fn), errors="replace") as f:
                    src = f.read()
                assert "oracle" not in src.lower(), f"{fn} mentions the oracle"


def test_argument_errors_need_no_gpu():
    bc = np.frombuffer(b"ACGTACGTACGTACGT" * 3, np.uint8).reshape(3, 16)
    for kwargs in (dict(max_error=16), dict(max_error=-1), dict(max_error=2, mem_gb=0.0), dict(max_error=2, threads=0)):
        with pytest.raises(_lib.SctError) as e:
            _lib.Matcher(bc, **kwargs)
        assert e.value.code == -1
    # -t so large that k >= bl - mr: the reference leaks its sentinel read id there -> rejected
    with pytest.raises(_lib.SctError):
        _lib.Matcher(bc[:, :4].copy(), 1, threads=256)


@pytest.mark.skipif(_lib.device_count() > 0, reason="only meaningful on a box without a GPU")
def test_no_cpu_fallback_without_gpu():
    bc = np.frombuffer(b"ACGTACGTACGTACGT", np.uint8).reshape(1, 16)
    with pytest.raises(_lib.SctError) as e:
        _lib.Matcher(bc, 2)
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)
    with pytest.raises(_lib.SctError):
        _lib.measure_int32_peak()


def test_prefix_k_matches_reference_rule():
    from sctagger_b200.match_trie import prefix_k
    # scTagger.py:689-697: smallest k with 4^k >= threads
    assert [prefix_k(t) for t in (1, 2, 4, 5, 16, 17, 64, 65, 256, 257)] == [0, 1, 1, 2, 2, 3, 3, 4, 4, 5]
