"""extract_sr_bc_from_lr (reference scTagger.py:792-844): exact whitelist search.

CPU: the oracle (oracle/wl_oracle.py) and the host mirror's selection / readers / writer against the golden outputs of
the UNMODIFIED reference (tests/golden/wl, made by tools/make_golden_wl.py with only `ahocorasick` substituted).
GPU: the CUDA hash-probe path (sct_wl_*) against the oracle per segment, and the whole CLI against the golden outputs."""
import gzip
import json
import os

import numpy as np
import pytest

import host_model as hm
from oracle import wl_oracle
from sctagger_b200 import _lib, cli, extract_sr_bc_from_lr as wl

D = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "wl")
CASES = json.load(open(os.path.join(D, "manifest.json")))["cases"]


def _read(path):
    with (gzip.open if path.endswith(".gz") else open)(path, "rb") as f:
        return f.read()


def _unpack(b, off):
    raw, o = np.asarray(b).tobytes(), np.asarray(off).tolist()
    return [raw[o[i]:o[i + 1]] for i in range(len(o) - 1)]


def _model_inputs(c):
    words = [l[:-1] for l in _read(os.path.join(D, c["wl"])).decode().splitlines(keepends=True)]
    segs = []
    for l in _read(os.path.join(D, c["lr"])).decode().splitlines():
        _, _, p, seg = l.split("\t")
        if p != "NA":
            segs.append(seg)
    return words, segs


def _args(c, out=None):
    argv = ["extract_sr_bc_from_lr", "-i", os.path.join(D, c["lr"]), "-wl", os.path.join(D, c["wl"])] + c["args"]
    return cli.parse_args(argv + (["-o", out] if out else []))


@pytest.mark.parametrize("c", CASES, ids=[c["out"] for c in CASES])
def test_oracle_and_host_selection_reproduce_the_reference(c):
    words, segs = _model_inputs(c)
    _, _, C = wl_oracle.search(words, segs)
    a = _args(c)
    want = _read(os.path.join(D, c["out"]))
    top = wl_oracle.select_top(C, a.thresh, a.step_size, a.max_barcode_cnt)
    assert "".join(f"{words[b]}\t{n}\n" for b, n in top).encode() == want
    idx = np.array(list(C.keys()), dtype=np.int32)
    cnt = np.array(list(C.values()), dtype=np.int64)
    wb, wo = hm.pack([w.encode() for w in words])
    assert wl.format_rows(wb, wo, *wl.select_top(idx, cnt, a.thresh, a.step_size, a.max_barcode_cnt)) == want


def test_select_top_with_nothing_counted_fails_like_the_reference():
    with pytest.raises(UnboundLocalError):
        wl.select_top(np.zeros(0, np.int32), np.zeros(0, np.int64), 0.005, 1000, 25000)
    with pytest.raises(UnboundLocalError):
        wl_oracle.select_top({}, 0.005, 1000, 25000)


@pytest.mark.parametrize("c", CASES[::4], ids=[c["wl"] for c in CASES[::4]])
def test_native_readers_match_the_reference_readers(c):
    words, segs = _model_inputs(c)
    wb, wo = wl.read_whitelist(os.path.join(D, c["wl"]), 3)
    assert _unpack(wb, wo) == [w.encode() for w in words]
    sb, so = wl.read_segments(os.path.join(D, c["lr"]), 3)
    assert _unpack(sb, so) == [s.encode() for s in segs]


@pytest.mark.parametrize("bad", [b"r\t1\t2\n", b"r\t1\t2\tACGT\textra\n"])
def test_segment_rows_need_exactly_four_columns(bad, tmp_path):
    p = tmp_path / "lr.tsv"
    p.write_bytes(b"ok\t0\t5\tACGT\n" + bad)
    with pytest.raises(ValueError):
        wl.read_segments(str(p), 2)


def test_numpy_twin_agrees_with_the_dictionary_oracle():
    rng = np.random.default_rng(3)
    words = ["".join("ACGT"[i] for i in rng.integers(0, 4, 12)) for _ in range(200)]
    words[7] = wl_oracle.rev_compl(words[1])
    segs = []
    for i in range(3000):
        w = words[int(rng.integers(0, 200))]
        pre = "".join("ACGTN"[i] for i in rng.integers(0, 5, int(rng.integers(0, 6))))
        segs.append(pre + (w if i % 3 else wl_oracle.rev_compl(w)) + pre[::-1] + (words[3] if i % 11 == 0 else ""))
    nh, first, _ = wl_oracle.search(words, segs)
    sb, so = hm.pack([s.encode() for s in segs])
    nh2, first2 = wl_oracle.search_numpy(words, sb, so)
    assert np.array_equal(nh, nh2)
    one = nh == 1
    assert np.array_equal(first[one], first2[one])


# ---- GPU -------------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("c", CASES[::4], ids=[c["wl"] for c in CASES[::4]])
def test_gpu_search_matches_the_oracle_per_segment(c):
    words, segs = _model_inputs(c)
    nh, first, C = wl_oracle.search(words, segs)
    wb, wo = hm.pack([w.encode() for w in words])
    sb, so = hm.pack([s.encode() for s in segs])
    with _lib.Whitelist(wb, wo) as t:
        got_nh, got_first = t.count(sb, so, per_segment=True)
        idx, cnt = t.result()
        assert np.array_equal(got_nh, nh)
        one = nh == 1
        assert np.array_equal(got_first[one], first[one])
        assert np.array_equal(got_first[nh == 0], first[nh == 0])
        assert idx.tolist() == list(C.keys()) and cnt.tolist() == list(C.values())
        # counts accumulate over calls and keep the job-wide first-seen order; reset clears them
        t.reset()
        half = len(segs) // 2
        t.count(*hm.pack([s.encode() for s in segs[:half]]))
        t.count(*hm.pack([s.encode() for s in segs[half:]]))
        idx2, cnt2 = t.result()
        assert np.array_equal(idx, idx2) and np.array_equal(cnt, cnt2)


@pytest.mark.gpu
@pytest.mark.parametrize("c", CASES, ids=[c["out"] for c in CASES])
def test_gpu_cli_reproduces_the_reference_output(c, tmp_path):
    out = str(tmp_path / "out.tsv.gz")
    argv = ["extract_sr_bc_from_lr", "-i", os.path.join(D, c["lr"]), "-wl", os.path.join(D, c["wl"]), "-o", out] + c["args"]
    assert cli.main(argv) == 0
    assert _read(out) == _read(os.path.join(D, c["out"]))


@pytest.mark.gpu
def test_gpu_large_random_whitelist_against_numpy_twin():
    """200 k 16-mers, 300 k segments through small chunks: n_hits everywhere, the word where there is exactly one."""
    rng = np.random.default_rng(11)
    W, n = 200_000, 300_000
    wm = rng.integers(0, 4, size=(W, 16))
    words = ["".join("ACGT"[i] for i in row) for row in wm[:2000]]
    lut = np.frombuffer(b"ACGT", np.uint8)
    wbytes = lut[wm].reshape(-1)
    woff = np.arange(W + 1, dtype=np.int64) * 16
    pick = rng.integers(0, W, n)
    rc = rng.random(n) < 0.5
    core = np.where(rc[:, None], 3 - wm[pick][:, ::-1], wm[pick])
    kind = rng.random(n)
    core[kind < 0.2, 7] ^= 1                              # one substitution: usually no hit
    pre, post = rng.integers(0, 4, (n, 3)), rng.integers(0, 4, (n, 4))
    seg = lut[np.concatenate([pre, core, post], axis=1)]
    seg[kind > 0.97, 1] = ord("N")
    sb = seg.reshape(-1)
    so = np.arange(n + 1, dtype=np.int64) * 23
    all_words = ["".join("ACGT"[i] for i in row) for row in wm]
    nh, first = wl_oracle.search_numpy(all_words, sb, so)
    with _lib.Whitelist(wbytes, woff) as t:
        got_nh, got_first = t.count(sb, so, per_segment=True)
        st = t.stats()
    assert np.array_equal(got_nh, nh)
    one = nh == 1
    assert one.sum() > 0.7 * n and np.array_equal(got_first[one], first[one])
    assert st["table_keys"] <= 2 * W and st["table_slots"] >= 4 * W and words


def test_semantics_published_by_pyahocorasick():
    """The example of pyahocorasick's own documentation: keys he / her / hers / she, values (insertion index, key),
        for item in A.iter("_hershe_"): ...  ->  (2, (0, 'he')) (3, (1, 'her')) (4, (2, 'hers')) (6, (3, 'she')) (6, (0, 'he'))
    — every occurrence is reported with its END index, overlapping ones and a key that is a suffix of another key included
    (the order of two hits with the same end index is irrelevant to the stage, which only counts hits).  And the documented
    dict-like add_word: a key added again keeps one entry and takes the new value (returns False)."""
    from oracle import wl_oracle
    A = wl_oracle.Automaton()
    for idx, key in enumerate("he her hers she".split()):
        assert A.add_word(key, (idx, key)) is True
    A.make_automaton()
    got = list(A.iter("_hershe_"))
    assert sorted(got) == sorted([(2, (0, "he")), (3, (1, "her")), (4, (2, "hers")), (6, (3, "she")), (6, (0, "he"))])
    assert [e for e, _ in got] == sorted(e for e, _ in got)
    assert A.add_word("she", (9, "SHE")) is False
    A.make_automaton()
    assert (6, (9, "SHE")) in list(A.iter("_hershe_")) and (6, (3, "she")) not in list(A.iter("_hershe_"))


@pytest.mark.gpu
def test_page_locked_arrays_and_compacted_result():
    """sct_wl_count copies from / to page-locked caller arrays directly (offsets in the caller's numbering, a sub-range
    included) and sct_wl_result reads back only the counted words: same per-segment hits and the same Counter (order
    included) as with pageable arrays, over several chunks (2.3 M segments > one chunk of 2 Mi) and several calls."""
    from sctagger_b200 import _lib
    rng = np.random.default_rng(12)
    acgt = np.frombuffer(b"ACGT", np.uint8)
    words = acgt[rng.integers(0, 4, size=(5000, 16))]
    n = 2_300_000
    segs = acgt[rng.integers(0, 4, size=(n, 20))]
    planted = np.nonzero(rng.random(n) < 0.6)[0]
    segs[planted, 2:18] = words[rng.integers(0, 3000, size=planted.size)]
    sb = segs.reshape(-1)
    so = np.arange(n + 1, dtype=np.int64) * 20
    wb, wo = words.reshape(-1), np.arange(5001, dtype=np.int64) * 16
    psb, pso = _lib.Pinned.copy_of(sb), _lib.Pinned.copy_of(so)
    pnh, phw = _lib.Pinned(n, np.int32), _lib.Pinned(n, np.int32)
    with _lib.Whitelist(wb, wo, threads=4) as a, _lib.Whitelist(wb, wo, threads=4) as b:
        nh, hw = a.count(sb, so, per_segment=True)
        nh2, hw2 = b.count(psb.array, pso.array, per_segment=True, out=(pnh.array, phw.array))
        assert b.stats()["chunks"] >= 2
        assert np.array_equal(nh, nh2) and np.array_equal(hw, hw2) and int((nh == 1).sum()) > n // 2
        for x, y in zip(a.result(), b.result()):
            assert np.array_equal(x, y)
        a.count(sb[so[1000]:so[3000]], so[1000:3001] - so[1000])      # a second call: the job-wide first-seen order goes on
        b.count(psb.array, pso.array[1000:3001])                      # the same range as a view of the pinned arrays
        ra, rb = a.result(), b.result()
        assert np.array_equal(ra[0], rb[0]) and np.array_equal(ra[1], rb[1]) and ra[0].size > 2000
    for x in (psb, pso, pnh, phw):
        x.close()
