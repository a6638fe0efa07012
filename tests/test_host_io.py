"""Native host I/O (sct_io_*: readers, writers, gzip) against the pure-Python restatement of the reference's readers
and writers (tests/host_model.py; reference scTagger.py:159-173, 602-628, 777-789, 317-320).  No GPU involved."""
import gzip
import os
import tempfile

import numpy as np
import pytest

from sctagger_b200 import _lib, match_trie as mt, extract_lr_bc as lr
import host_model as hm


def _unpack(b, off):
    raw, o = np.asarray(b).tobytes(), np.asarray(off).tolist()
    return [raw[o[i]:o[i + 1]] for i in range(len(o) - 1)]


def _write(path, data, gz):
    if gz:
        with gzip.open(path, "wb") as f:
            f.write(data)
    else:
        with open(path, "wb") as f:
            f.write(data)


def _random_tsv(rng, n, crlf=False, trailing_newline=True):
    rows = []
    for i in range(n):
        name = b"read%d/%d" % (i, rng.integers(0, 10 ** 6))
        seg = bytes(rng.choice(np.frombuffer(b"ACGTNacgt", np.uint8), size=int(rng.integers(0, 40))))
        cols = [name, b"%d" % rng.integers(-1, 5), b"%d" % rng.integers(-50, 50), seg]
        cols += [b"extra"] * int(rng.integers(0, 3))
        rows.append(b"\t".join(cols))
    eol = b"\r\n" if crlf else b"\n"
    data = eol.join(rows)
    if trailing_newline and rows:
        data += eol
    return data


@pytest.mark.parametrize("gz", [False, True])
@pytest.mark.parametrize("n,crlf,trail", [(0, False, True), (1, False, False), (7, True, True), (5000, False, True),
                                           (40000, False, False)])
@pytest.mark.parametrize("threads", [1, 5])
def test_read_long_reads_matches_model(gz, n, crlf, trail, threads):
    rng = np.random.default_rng(n + 17 * gz + threads)
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "lr.tsv" + (".gz" if gz else ""))
        _write(p, _random_tsv(rng, n, crlf, trail), gz)
        names, segs = hm.read_long_reads(p)
        got = mt.read_long_reads(p, threads)
        assert len(got) == n
        assert _unpack(got.name_bytes, got.name_offsets) == names
        assert _unpack(got.seg_bytes, got.offsets) == segs
        assert got.names == names


@pytest.mark.parametrize("bad", [b"a\tb\tc\n", b"ok\t1\t2\tACGT\n\nok2\t1\t2\tACGT\n", b"ok\t1\t2\tACGT\nshort\t1\n"])
def test_read_long_reads_short_row_is_index_error(bad):
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "lr.tsv")
        _write(p, bad, False)
        with pytest.raises(IndexError):
            hm.read_long_reads(p)
        with pytest.raises(IndexError):
            mt.read_long_reads(p, 3)


def test_read_errors_are_os_errors():
    with tempfile.TemporaryDirectory() as d:
        with pytest.raises(OSError):
            mt.read_long_reads(os.path.join(d, "missing.tsv"))
        p = os.path.join(d, "plain.tsv.gz")
        _write(p, b"r\t1\t2\tACGT\n", False)         # named .gz, is not gzip: gzip.open raises BadGzipFile (an OSError)
        with pytest.raises(OSError):
            mt.read_long_reads(p)


def test_read_multi_member_gzip():
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "lr.tsv.gz")
        with open(p, "wb") as f:
            f.write(gzip.compress(b"r0\t0\t0\tACGT\nr1\t0\t0\t") + gzip.compress(b"GGGG\nr2\t0\t0\tTT\n"))
        got = mt.read_long_reads(p, 2)
        assert got.names == [b"r0", b"r1", b"r2"]
        assert _unpack(got.seg_bytes, got.offsets) == [b"ACGT", b"GGGG", b"TT"]


@pytest.mark.parametrize("gz", [False, True])
@pytest.mark.parametrize("n,tail", [(0, b""), (3, b""), (2500, b""), (9, b"@dangling header\n"), (4, b"@h\nACGT  \t\n+")])
def test_read_fastqs_matches_model(gz, n, tail):
    rng = np.random.default_rng(n)
    recs = []
    for i in range(n):
        seq = bytes(rng.choice(np.frombuffer(b"ACGTN", np.uint8), size=int(rng.integers(0, 300))))
        recs.append(b"@r%d some comment\t%d\n%s%s\n+\n%s\n" % (i, i, seq, b" \r" if i % 7 == 0 else b"", b"I" * len(seq)))
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "reads.fastq" + (".gz" if gz else ""))
        _write(p, b"".join(recs) + tail, gz)
        names, seqs = hm.read_fastq(p)
        got = lr.read_fastqs([p], False, 4)
        assert _unpack(got.name_bytes, got.name_offsets) == names
        assert _unpack(got.read_bytes, got.offsets) == seqs
        two = lr.read_fastqs([p, p], False, 2)
        assert two.names == names + names and _unpack(two.read_bytes, two.offsets) == seqs + seqs


def _random_result(rng, n, n_barcodes):
    min_ed = rng.integers(-1, 4, size=n).astype(np.int32)
    cnt = np.where(min_ed >= 0, rng.integers(1, 5, size=n), 0)
    cnt[rng.random(n) < 0.01] *= 30
    cnt = np.where(min_ed >= 0, np.maximum(cnt, 1), 0)
    off = np.zeros(n + 1, dtype=np.int64)
    off[1:] = np.cumsum(cnt)
    ids = rng.integers(0, 2 * n_barcodes, size=int(off[-1])).astype(np.int32)
    return mt.MatchResult(min_ed, cnt.astype(np.int32), off, ids)


@pytest.mark.parametrize("n", [0, 1, 300, 50000])
@pytest.mark.parametrize("level", [0, 1, 6])
def test_format_matches_matches_model(n, level):
    rng = np.random.default_rng(n + level)
    barcodes = [bytes(rng.choice(np.frombuffer(b"ACGTN", np.uint8), size=16)) for _ in range(50)]
    names = [b"name_%d" % i for i in range(n)]
    segs = [bytes(rng.choice(np.frombuffer(b"ACGTn", np.uint8), size=int(rng.integers(0, 30)))) for _ in range(n)]
    res = _random_result(rng, n, len(barcodes))
    want = hm.format_matches(names, segs, res.min_ed, res.n_ties, res.tie_offsets, res.tie_ids, barcodes)
    nb, no = hm.pack(names)
    sb, so = hm.pack(segs)
    reads = mt.LongReads(nb, no, sb, so)
    got = mt.format_reads_rows(reads, res, [b.decode() for b in barcodes], gzip_level=level, threads=3)
    if level:
        assert got[:2] == b"\x1f\x8b"
        got = gzip.decompress(got)
    assert got == want


def test_format_matches_rejects_bad_tie_id():
    res = mt.MatchResult(np.array([0], np.int32), np.array([1], np.int32), np.array([0, 1], np.int64), np.array([999], np.int32))
    nb, no = hm.pack([b"r"])
    sb, so = hm.pack([b"ACGT"])
    with pytest.raises(_lib.SctError):
        mt.format_reads_rows(mt.LongReads(nb, no, sb, so), res, ["ACGTACGTACGTACGT"])


@pytest.mark.parametrize("n", [0, 5, 6000])
@pytest.mark.parametrize("level", [0, 6])
def test_format_lr_bc_matches_model(n, level):
    rng = np.random.default_rng(n + level)
    names = [b"r%d" % i for i in range(n)]
    seqs = [bytes(rng.choice(np.frombuffer(b"ACGT", np.uint8), size=int(rng.integers(0, 120)))) for _ in range(n)]
    dist = rng.integers(-1, 6, size=n).astype(np.int64)
    s = rng.integers(-150, 150, size=n).astype(np.int64)
    e = rng.integers(-150, 150, size=n).astype(np.int64)
    e[rng.random(n) < 0.2] = 0
    loc = rng.integers(-150, 150, size=n).astype(np.int64)
    want = hm.format_lr_bc(names, seqs, dist, loc, s, e)
    nb, no = hm.pack(names)
    rb, ro = hm.pack(seqs)
    got = lr.format_rows(lr.Reads(nb, no, rb, ro), (dist, loc, s, e), gzip_level=level, threads=3)
    if level:
        got = gzip.decompress(got)
    assert got == want


def test_match_trie_writes_gzip_members_the_reference_reader_accepts(tmp_path):
    """A multi-member .gz written by the native writer reads back through gzip.open (what the next stage of the
    reference pipeline uses) and through the native reader."""
    rng = np.random.default_rng(5)
    n = 40000
    names = [b"r%d" % i for i in range(n)]
    segs = [b"ACGTACGTACGTACGTACGTAC"] * n
    res = _random_result(rng, n, 10)
    nb, no = hm.pack(names)
    sb, so = hm.pack(segs)
    bcs = ["ACGTACGTACGTACGA"] * 10
    plain = mt.format_reads_rows(mt.LongReads(nb, no, sb, so), res, bcs)
    gz = mt.format_reads_rows(mt.LongReads(nb, no, sb, so), res, bcs, gzip_level=6)
    p = tmp_path / "out.tsv.gz"
    p.write_bytes(gz)
    with gzip.open(p, "rb") as f:
        assert f.read() == plain
    back = mt.read_long_reads(str(p), 4)          # 5 columns: name, e, n, seg, barcodes -> column 3 = the segment
    assert len(back) == int((res.min_ed >= 0).sum())


def test_indexed_gzip_members_parallel_and_mixed_files(tmp_path):
    """The native writer's gzip members carry their sizes in the header's extra field ("SC"): a file made only of such
    members is inflated in parallel; as soon as one member is foreign (the reference's writer, plain `gzip`) or a size is
    wrong, the reader takes the serial zlib path — same rows either way."""
    n = 50000
    names = [b"r%d" % i for i in range(n)]
    seqs = [b"ACGTACGTACGTACGTACGTAC"] * n
    nb, no = hm.pack(names)
    rb, ro = hm.pack(seqs)
    z = np.zeros(n, np.int64)
    plain = lr.format_rows(lr.Reads(nb, no, rb, ro), (z, z, z, z))
    ours = lr.format_rows(lr.Reads(nb, no, rb, ro), (z, z, z, z), gzip_level=1, threads=4)
    assert ours[3] & 4 and ours[12:14] == b"SC"                       # FEXTRA with the size subfield
    csize = int.from_bytes(ours[16:20], "little")
    assert ours[csize:csize + 2] == b"\x1f\x8b" and gzip.decompress(ours) == plain
    foreign = gzip.compress(b"tail\t0\t0\tGGGG\n")
    broken = bytearray(ours)
    broken[20] ^= 0x01                                                # wrong uncompressed size in the first member
    for name, data, extra in (("ours.tsv.gz", ours, 0), ("mixed.tsv.gz", ours + foreign, 1),
                              ("mixed2.tsv.gz", foreign + ours, 1), ("broken.tsv.gz", bytes(broken), 0)):
        p = tmp_path / name
        p.write_bytes(data)
        got = mt.read_long_reads(str(p), 4)
        assert len(got) == n + extra
        segs = _unpack(got.seg_bytes, got.offsets)
        assert segs.count(b"ACGTACGTACGTACGTACGTAC") == n and (extra == 0 or b"GGGG" in segs)


@pytest.mark.parametrize("cut", [1, 9, 30, "half"])
def test_truncated_or_corrupt_gzip_is_an_os_error(tmp_path, cut):
    """gzip.open raises EOFError / BadGzipFile on a truncated or corrupt stream (the reference's readers,
    scTagger.py:605, 634); the native reader must not hand back the rows inflated so far."""
    rows = b"".join(b"r%d\t0\t0\tACGTACGTACGTACGTACGTAC\n" % i for i in range(20000))
    whole = gzip.compress(rows)
    data = whole[: len(whole) // 2] if cut == "half" else whole[:-cut]
    p = tmp_path / "trunc.tsv.gz"
    p.write_bytes(data)
    with pytest.raises(EOFError):
        gzip.open(p, "rb").read()
    with pytest.raises(OSError):
        mt.read_long_reads(str(p), 4)
    # the native writer's multi-member files: a truncated last member must fail as well, not fall back silently
    names = [b"r%d" % i for i in range(40000)]
    nb, no = hm.pack(names)
    rb, ro = hm.pack([b"ACGTACGTACGTACGTACGTAC"] * len(names))
    z = np.zeros(len(names), np.int64)
    ours = lr.format_rows(lr.Reads(nb, no, rb, ro), (z, z, z, z), gzip_level=1, threads=4)
    q = tmp_path / "trunc_members.tsv.gz"
    q.write_bytes(ours[: len(ours) // 2] if cut == "half" else ours[:-cut])
    with pytest.raises(OSError):
        mt.read_long_reads(str(q), 4)
    bad = bytearray(whole)
    bad[len(bad) // 2] ^= 0xFF                                            # corrupt deflate data / CRC mismatch
    p.write_bytes(bytes(bad))
    with pytest.raises(OSError):
        mt.read_long_reads(str(p), 4)
