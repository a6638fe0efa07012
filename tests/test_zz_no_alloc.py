"""GPU: the allocation-free entry point sct_align_strand_dist_noalloc (static device scratch, no cudaMalloc) against the
existing entry point sct_align, through the C ABI, on two calls in a row.  Integer results: bit-exact.
Named test_zz_* so that pytest collects it last."""
import numpy as np
import pytest

from sctagger_b200 import _lib, synth

pytestmark = pytest.mark.gpu


def _reference(adapter, rb, off):
    with _lib.Aligner(adapter, threads=4) as al:
        strand, dist, _, _ = al.align(rb, off)
    return np.array(strand), np.array(dist)


def test_noalloc_entry_point_equals_sct_align_on_two_calls_in_a_row():
    # call 1: the config-4 generator (both strands, reads without adapter); call 2: fewer and shorter reads, another
    # adapter (m = 15), non-ACGT bytes — the shared scratch still holds call 1's contents behind them
    rb1, off1 = synth.make_reads(1500, 4242)
    rb2, off2 = synth.make_reads(400, 99, adapter=b"ACGTACGTACGTAAC", median_len=300, min_len=1, max_len=900)
    rb2 = rb2.copy()
    rb2[::97] = ord("N")
    for adapter, rb, off in ((synth.ADAPTER, rb1, off1), (b"ACGTACGTACGTAAC", rb2, off2), (synth.ADAPTER, rb1, off1)):
        want_s, want_d = _reference(adapter, rb, off)
        got_s, got_d = _lib.align_strand_dist_noalloc(adapter, rb, off)
        assert np.array_equal(got_s, want_s)
        assert np.array_equal(got_d, want_d)
        assert (got_d >= 0).sum() > 100
    # a slice of a larger input: offsets in the caller's numbering (read_offsets[0] != 0)
    got_s, got_d = _lib.align_strand_dist_noalloc(synth.ADAPTER, rb1, off1[200:701])
    want_s, want_d = _reference(synth.ADAPTER, rb1, off1)
    assert np.array_equal(got_s, want_s[200:700]) and np.array_equal(got_d, want_d[200:700])


def test_noalloc_entry_point_rejects_sizes_beyond_its_scratch():
    one = np.frombuffer(b"ACGT" * 8, np.uint8)
    too_many = np.arange(8194, dtype=np.int64)                    # 8193 reads of one base
    with pytest.raises(_lib.SctError) as e:
        _lib.align_strand_dist_noalloc(synth.ADAPTER, np.tile(one, 300), too_many)
    assert e.value.code == -4 and "8192" in str(e.value)
    big = np.zeros((2 << 20) + 16, np.uint8) + ord("A")
    with pytest.raises(_lib.SctError) as e:                        # 2 MiB + 16 bytes in reads of 8 kb
        _lib.align_strand_dist_noalloc(synth.ADAPTER, big, np.append(np.arange(0, big.size, 8192), big.size).astype(np.int64))
    assert e.value.code == -4
    with pytest.raises(_lib.SctError) as e:                        # one read of 20 000 bases
        _lib.align_strand_dist_noalloc(synth.ADAPTER, big[:20000], np.array([0, 20000], np.int64))
    assert e.value.code == -4
    s, d = _lib.align_strand_dist_noalloc(synth.ADAPTER, one, np.zeros(1, np.int64))     # n = 0
    assert s.size == 0 and d.size == 0
