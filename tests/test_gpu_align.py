"""GPU parity of the extract_lr_bc adapter alignment (sct_align, through the C ABI) against the CPU oracle
(oracle/align_oracle.c, edlib HW semantics restated — parity with edlib itself is unpinned).  Bit-exact: integer work."""
import numpy as np
import pytest

from oracle import oracle
from sctagger_b200 import _lib, synth

pytestmark = pytest.mark.gpu


def _pack(reads):
    off = np.zeros(len(reads) + 1, np.int64)
    np.cumsum([len(r) for r in reads], out=off[1:])
    return np.frombuffer(b"".join(reads), np.uint8), off


def _same(got, want):
    for g, w, name in zip(got, want, ("strand", "dist", "loc_offsets", "locs")):
        assert np.array_equal(g, w), name


def test_hand_cases_and_edges():
    a1 = b"ACGTT"
    reads = [b"GGACGTTGG", b"GGAACGTGG", b"", b"GGGGGG", b"ACGTTACGTT", b"TTACGTCC", b"A", b"ACG", b"N", b"acgtt",
             b"ACGTNACGTT" * 30, b"AACGT" * 50, b"T" * 17, b"AACGTACGTT"]
    rb, off = _pack(reads)
    _same(_lib.align_reads(a1, rb, off), oracle.align(a1, rb, off))
    # n = 0
    s, d, lo, locs = _lib.align_reads(a1, np.zeros(0, np.uint8), np.zeros(1, np.int64))
    assert s.size == 0 and lo.tolist() == [0] and locs.size == 0


@pytest.mark.parametrize("seed,adapter", [(1, synth.ADAPTER), (2, b"ACGTACGTAC"), (3, b"GATTACAGATTACAGATTACAGATTACAGATT"),
                                          (4, b"ANCGT"), (5, b"T"), (6, b"CTACACGACGCTCTTCCGATCTA")])
def test_random_reads_vs_oracle(seed, adapter):
    rb, off = synth.make_reads(3000, seed, adapter=adapter if (b"N" not in adapter and len(adapter) > 4) else synth.ADAPTER,
                               median_len=300, min_len=1, max_len=1500)
    rb = rb.copy()
    rng = np.random.default_rng(seed)
    rb[rng.integers(0, rb.size, size=500)] = ord("N")
    rb[rng.integers(0, rb.size, size=500)] = ord("a")
    want = oracle.align(adapter, rb, off)
    _same(_lib.align_reads(adapter, rb, off), want)
    # ragged chunking: by bytes and by reads
    _same(_lib.align_reads(adapter, rb, off, chunk_bytes=5000), want)
    _same(_lib.align_reads(adapter, rb, off, chunk_reads=7, chunk_bytes=1 << 20), want)


def test_cfg5_shape_sample_and_staged_timing():
    """BASELINE config 5 shape (log-normal lengths, median 900, 22-nt adapter) on a sample the DP oracle finishes fast."""
    rb, off = synth.make_reads(20_000, synth.SEEDS["cfg5"])
    want = oracle.align(synth.ADAPTER, rb, off, algo="bv")
    with _lib.Aligner(synth.ADAPTER, threads=8) as a:
        _same(a.align(rb, off), want)
        st = a.stats()
        assert st["reads"] == 20_000 and st["launches"] >= 8 and st["h2d_bytes"] >= rb.size
        a.stage(rb, off)
        ms, k1 = a.run_staged()
        assert ms > 0 and 0 < k1 <= ms


def test_long_reads_and_length_sorting():
    rng = np.random.default_rng(3)
    lens = [0, 1, 15, 16, 17, 31, 32, 33, 40_000, 100_000, 5, 70_000]
    reads = [bytes(synth.ACGT[rng.integers(0, 4, size=n)]) for n in lens]
    reads[8] = reads[8][:20_000] + synth.ADAPTER + reads[8][20_022:]
    rb, off = _pack(reads)
    _same(_lib.align_reads(synth.ADAPTER, rb, off), oracle.align(synth.ADAPTER, rb, off, algo="bv"))


def test_long_location_spans_take_the_warp_path():
    """Reads whose locations are far apart (no adapter at all; the adapter twice, 3 kb apart; a long homopolymer where
    every position ties) go through aln_pass2_long_kernel (one warp per read, span split over the lanes); reads with a
    span of a few columns stay on the thread-per-read kernel.  Both against the oracle, both strands."""
    rng = np.random.default_rng(21)
    lut = np.frombuffer(b"ACGT", np.uint8)
    a1 = synth.ADAPTER
    rc = bytes({65: 84, 67: 71, 71: 67, 84: 65}[c] for c in reversed(a1))
    rand = lambda n: lut[rng.integers(0, 4, n)].tobytes()
    reads = []
    for i in range(200):
        k = i % 5
        if k == 0:
            reads.append(rand(int(rng.integers(500, 5000))))                              # no adapter
        elif k == 1:
            reads.append(rand(40) + a1 + rand(3000) + a1 + rand(int(rng.integers(0, 50))))   # '+' twice, far apart
        elif k == 2:
            reads.append(rand(30) + rc + rand(2500) + rc + rand(20))                      # '-' twice: two start searches
        elif k == 3:
            reads.append(b"T" * int(rng.integers(400, 3000)))                             # every column ties
        else:
            reads.append(rand(20) + a1 + rand(int(rng.integers(100, 2000))))              # short span
    rb, off = _pack(reads)
    got = _lib.align_reads(a1, rb, off)
    _same(got, oracle.align(a1, rb, off))
    assert int(np.diff(got[2]).max()) > 300                                               # the homopolymer reads


def test_very_long_reads_take_the_warp_per_read_pass1():
    """Reads longer than 16 384 bases are scored by aln_pass1_long_kernel (lane l scans the l-th 1/32 of the read behind a
    warm-up): adapters at piece seams, at both ends, on both strands, twice, not at all, lengths around the threshold and
    around the length-bucket boundary, mixed with ordinary reads — against the oracle."""
    rng = np.random.default_rng(33)
    lut = np.frombuffer(b"ACGT", np.uint8)
    a1 = synth.ADAPTER
    rc = bytes({65: 84, 67: 71, 71: 67, 84: 65}[c] for c in reversed(a1))
    rand = lambda n: lut[rng.integers(0, 4, n)].tobytes()

    def plant(n, what, at):
        b = bytearray(rand(n))
        for w, a in zip(what, at):
            b[a:a + len(w)] = w
        return bytes(b)

    reads = [rand(16384), rand(16385), rand(16415), rand(16416), rand(32735), rand(32736), rand(40000)]
    for n in (20000, 50001, 120000):
        piece = (n + 31) // 32
        reads += [plant(n, [a1], [5]), plant(n, [rc], [n - 30]), plant(n, [a1], [piece * 7 - 11]),     # across a piece seam
                  plant(n, [rc], [piece * 19 - 3]), plant(n, [a1, a1], [piece * 3, piece * 20 + 5]),    # twice, far apart
                  plant(n, [a1[:-3]], [piece * 11 - 9]), plant(n, [rc, a1[:-2]], [100, n - 500])]
    reads += [rand(int(rng.integers(200, 3000))) for _ in range(40)] + [plant(900, [a1], [33]), plant(900, [rc], [850])]
    order = rng.permutation(len(reads))
    reads = [reads[i] for i in order]
    rb, off = _pack(reads)
    want = oracle.align(a1, rb, off, workers=8, algo="bv")
    _same(_lib.align_reads(a1, rb, off), want)
    _same(_lib.align_reads(b"ACGTACGTACGTAAC", rb, off), oracle.align(b"ACGTACGTACGTAAC", rb, off, workers=8, algo="bv"))   # m = 15: generic pass 1


def test_align_errors():
    with pytest.raises(_lib.SctError) as e:
        _lib.Aligner(b"A" * 33)
    assert e.value.code == -4
    with pytest.raises(_lib.SctError):
        _lib.Aligner(b"")
    with pytest.raises(_lib.SctError):
        _lib.Aligner(synth.ADAPTER, threads=0)
    with _lib.Aligner(synth.ADAPTER) as a:
        with pytest.raises(_lib.SctError):
            a.align(np.zeros(4, np.uint8), np.array([0, 3, 2], np.int64))
        with pytest.raises(_lib.SctError):
            a.run_staged()


def test_page_locked_buffers_are_used_directly_and_bad_offsets_fail():
    """sct_align copies from / to page-locked caller arrays without a staging pass: same result as with pageable arrays,
    over several chunks; offsets that decrease fail the call (validated on the device)."""
    rb, off = synth.make_reads(3000, 77)
    with _lib.Aligner(synth.ADAPTER, threads=4, chunk_reads=700) as al:
        want = al.align(rb, off)
        prb, poff, out = _lib.Pinned.copy_of(rb), _lib.Pinned.copy_of(off), _lib.AlignBuffers(3000)
        got = al.align(prb.array, poff.array, out=out, copy_locs=False)
        assert al.stats()["chunks"] >= 4
        for x, y in zip(got, want):
            assert np.array_equal(x, y)
        bad = off.copy()
        bad[10] = bad[12] + 5
        with pytest.raises(_lib.SctError):
            al.align(rb, bad)
        again = al.align(rb, off)                       # the handle stays usable
        for x, y in zip(again, want):
            assert np.array_equal(x, y)
        for x in (prb, poff, out):
            x.close()
