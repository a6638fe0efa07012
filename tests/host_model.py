"""Pure-Python restatements of the reference's readers and writers (reference scTagger.py:159-173, 602-628, 777-789,
317-320) — test infrastructure: the native sct_io_* functions are checked against these on random and edge-case files."""
import gzip

import numpy as np

_RC = {ord("A"): ord("T"), ord("C"): ord("G"), ord("G"): ord("C"), ord("T"): ord("A")}


def rev_compl(b: bytes) -> bytes:
    return bytes(_RC.get(c, c) for c in reversed(b))


def _read_all(path, gz):
    if gz:
        with gzip.open(path, "rb") as f:
            return f.read()
    with open(path, "rb") as f:
        return f.read()


def _lines(data):
    lines = data.split(b"\n")
    if lines and lines[-1] == b"":
        lines.pop()
    return lines


def read_long_reads(path):
    """-> (names list[bytes], segs list[bytes]); IndexError on rows with fewer than four columns."""
    names, segs = [], []
    for l in _lines(_read_all(path, path.endswith(".gz"))):
        cols = l.rstrip(b"\r").split(b"\t")
        names.append(cols[0])
        segs.append(cols[3])
    return names, segs


def read_fastq(path, gzipped=False):
    lines = _lines(_read_all(path, gzipped or path.endswith(".gz")))
    names = [l.split()[0][1:] for l in lines[0::4]]
    seqs = [l.rstrip() for l in lines[1::4]]
    n = min(len(names), len(seqs))
    return names[:n], seqs[:n]


def format_matches(names, segs, min_ed, n_ties, tie_offsets, tie_ids, barcodes):
    out = []
    for rid in range(len(names)):
        if min_ed[rid] < 0:
            continue
        parts = []
        for pid in tie_ids[tie_offsets[rid]:tie_offsets[rid + 1]]:
            b = barcodes[int(pid) >> 1]
            parts.append(b if int(pid) & 1 else rev_compl(b))
        out.append(b"%s\t%d\t%d\t%s\t%s\n" % (names[rid], min_ed[rid], n_ties[rid], segs[rid], b",".join(parts)))
    return b"".join(out)


def format_lr_bc(names, seqs, dist, loc, s, e):
    out = []
    for i in range(len(names)):
        if dist[i] < 0:
            out.append(b"%s\t-1\tNA\t\n" % names[i])
        else:
            out.append(b"%s\t%d\t%d\t%s\n" % (names[i], dist[i], loc[i], seqs[i][int(s[i]):int(e[i]) or None]))
    return b"".join(out)


def pack(items):
    off = np.zeros(len(items) + 1, dtype=np.int64)
    if items:
        off[1:] = np.cumsum([len(x) for x in items])
    return np.frombuffer(b"".join(items), dtype=np.uint8), off
