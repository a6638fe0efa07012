"""extract_lr_bc host mirror (sctagger_b200/extract_lr_bc.py) against outputs of the UNMODIFIED reference
(tests/golden/lr_bc, made by tools/make_golden_lr.py: the reference's own read_fastqs / get_alns / get_ranges /
filter_aln / output_matches with only `edlib` replaced by a restatement — edlib parity itself is unpinned).

CPU: range detection, filtering and formatting fed with the CPU oracle's alignments reproduce the reference's bytes.
GPU: the whole command line (`python -m sctagger_b200 extract_lr_bc ...`) reproduces them through the CUDA library."""
import contextlib
import gzip
import io
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import util
from oracle import oracle
from sctagger_b200 import cli
from sctagger_b200 import extract_lr_bc as lr

D = os.path.join(util.GOLD, "lr_bc")
with open(os.path.join(D, "manifest.json")) as _f:
    CASES = json.load(_f)["cases"]


def _argv(c, out=None):
    a = ["extract_lr_bc", "-r"] + [os.path.join(D, r) for r in c["reads"]] + c["args"]
    return a + (["-o", out] if out else [])


def _expected(c):
    with gzip.open(os.path.join(D, c["out"]), "rb") as f:
        return f.read()


@pytest.mark.parametrize("c", CASES, ids=lambda c: c["name"])
def test_host_logic_with_oracle_alignments(c):
    args = cli.parse_args(_argv(c))
    with contextlib.redirect_stderr(io.StringIO()):
        reads = lr.read_fastqs(args.reads, args.gzipped)
        alns = lr.Alignments(*oracle.align(args.short_read_adapter.encode(), reads.read_bytes, reads.offsets))
        if len(args.ranges[0]) + len(args.ranges[1]) == 0:
            rd = lr.set_global_range_dicts(lr.get_possible_ranges(alns))
        else:
            rd = lr.RangeDicts(args.ranges, inclusive_end=True)
        got = lr.format_rows(reads, lr.get_filter_alns(alns, rd, args.num_bp_after, args.threads))
    assert got == _expected(c)


def test_read_fastqs_and_tuples_view():
    with contextlib.redirect_stderr(io.StringIO()):
        reads = lr.read_fastqs([os.path.join(D, "readsA.fastq"), os.path.join(D, "readsB.fastq.gz")])
    assert len(reads) == 260 and reads.names[0] == b"r0" and reads.names[259] == b"r259"
    assert reads.seq(5) == b"" and reads.seq(6) == b"ACGT"
    a = lr.Alignments(*oracle.align(lr.DEFAULT_ADAPTER.encode(), reads.read_bytes, reads.offsets))
    t = a.tuples()
    assert t[5] == ("NA", -1, ["NA"]) and t[7] == ("+", 0, [21]) and t[9][:2] == ("+", 0) and t[9][2] == [21, 53]


def test_get_ranges_quirks():
    with contextlib.redirect_stderr(io.StringIO()):
        assert lr.get_ranges([]) == []
        # one dominant peak; bin 0 is never joined to a peak (the reference tests i - 1 > 0)
        data = [50] * 500 + [51] * 300 + [49] * 200 + [10] * 3
        assert lr.get_ranges(data) == [(49, 51)]
        data = [0] * 100 + [1] * 100 + [2] * 150
        assert lr.get_ranges(data)[0] == (1, 2)


def test_flag_parsing_matches_reference_rules():
    a = cli.parse_args(["extract_lr_bc", "-r", "x.fastq", "-g", "f20:40", "r1:30"])
    assert a.ranges == [[(19, 40)], [(-30, 0)]] and a.threads == 1 and a.num_bp_after == 20
    assert a.short_read_adapter == "CTACACGACGCTCTTCCGATCT" and a.gzipped is False and a.outfile is None
    for bad in (["-g", "x1:2"], ["-g", "f0:3"], ["-g", "f5:3"], ["-g", "f1"], ["-t", "0"], ["--num-bp-after", "0"]):
        with pytest.raises(AssertionError):
            cli.parse_args(["extract_lr_bc", "-r", "x.fastq"] + bad)
    with pytest.raises(AssertionError):      # overlapping ranges (scTagger.py:121-123)
        lr.RangeDicts(cli.parse_args(["extract_lr_bc", "-r", "x", "-g", "f1:10", "f10:20"]).ranges, inclusive_end=True)


@pytest.mark.gpu
@pytest.mark.parametrize("c", CASES, ids=lambda c: c["name"])
def test_cli_end_to_end_on_gpu(c, tmp_path):
    out = str(tmp_path / "out.tsv.gz")
    r = subprocess.run([sys.executable, "-m", "sctagger_b200"] + _argv(c, out), cwd=util.ROOT, capture_output=True)
    assert r.returncode == 0, r.stderr.decode()
    with gzip.open(out, "rb") as f:
        assert f.read() == _expected(c)


@pytest.mark.gpu
def test_cli_stdout_on_gpu():
    c = CASES[1]
    r = subprocess.run([sys.executable, "-m", "sctagger_b200"] + _argv(c), cwd=util.ROOT, capture_output=True)
    assert r.returncode == 0 and r.stdout == _expected(c)
