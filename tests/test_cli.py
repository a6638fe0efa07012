"""The `match_trie` command line (reference scTagger.py:81-97, 140-143): same flags, defaults and assertions; on a GPU
the whole CLI (`-lr/-sr -> -o`) reproduces the reference's output files byte for byte after decompression."""
import gzip
import os
import subprocess
import sys

import pytest

import util
from sctagger_b200 import cli


def test_defaults_match_the_reference():
    a = cli.parse_args(["match_trie", "-lr", "L", "-sr", "S"])
    assert (a.long_read_segments, a.short_read_barcodes) == ("L", "S")
    assert (a.max_error, a.mem, a.barcode_length, a.threads, a.plotfile, a.outfile) == (2, 16.0, 16, 16, None, None)
    a = cli.parse_args(["match_trie", "--long-read-segments", "L", "--short-read-barcodes", "S", "--max-error", "3",
                        "--mem", "2.5", "--barcode-length", "12", "--threads", "4", "--plotfile", "p.png",
                        "--outfile", "o.gz"])
    assert (a.max_error, a.mem, a.barcode_length, a.threads, a.plotfile, a.outfile) == (3, 2.5, 12, 4, "p.png", "o.gz")


@pytest.mark.parametrize("argv", [
    ["-m", "0"],                 # assert args.mem > 0                      scTagger.py:141
    ["-bl", "0"],                # assert args.barcode_length > 0           scTagger.py:142
    ["-bl", "2", "-mr", "2"],    # assert barcode_length > max_error >= 0   scTagger.py:143
    ["-mr", "-1"],
])
def test_assertions_match_the_reference(argv):
    with pytest.raises(AssertionError):
        cli.parse_args(["match_trie", "-lr", "L", "-sr", "S"] + argv)


def test_required_flags_and_unknown_subcommand():
    for argv in (["match_trie", "-lr", "L"], ["match_trie", "-sr", "S"], [], ["extract_sr_bc"]):
        with pytest.raises(SystemExit):
            cli.parse_args(argv)


def test_device_flags():
    assert cli.parse_args(["match_trie", "-lr", "L", "-sr", "S", "--gpus", "4"]).devices == [0, 1, 2, 3]
    assert cli.parse_args(["match_trie", "-lr", "L", "-sr", "S", "--devices", "2,5"]).devices == [2, 5]
    assert cli.parse_args(["match_trie", "-lr", "L", "-sr", "S"]).devices is None


@pytest.mark.gpu
@pytest.mark.parametrize("c", [c for c in util.golden_cases() if c["t"] in (16, 17)][:8], ids=util.case_id)
def test_cli_end_to_end_against_reference_output(c, tmp_path):
    d = os.path.join(util.GOLD, c["dir"])
    for out in ("out.tsv.gz", "out.tsv"):
        o = str(tmp_path / out)
        r = subprocess.run([sys.executable, "-m", "sctagger_b200", "match_trie", "-lr", os.path.join(d, c["lr"]),
                            "-sr", os.path.join(d, c["sr"]), "-mr", str(c["mr"]), "-bl", str(c["bl"]),
                            "-t", str(c["t"]), "-o", o], cwd=util.ROOT, capture_output=True)
        assert r.returncode == 0, r.stderr.decode()
        assert r.stdout == b""
        with open(o, "rb") as f:
            raw = f.read()
        assert (raw[:2] == b"\x1f\x8b") == out.endswith("gz")        # gzip iff the name ends with 'gz' (scTagger.py:771)
        got = gzip.decompress(raw) if out.endswith("gz") else raw
        assert got == util.read_maybe_gz(os.path.join(d, c["out"]))


@pytest.mark.gpu
def test_cli_stdout_and_barcode_length_assertion(tmp_path):
    c = [c for c in util.golden_cases() if c["dir"] == "edge16" and c["mr"] == 2 and c["t"] == 16][0]
    d = os.path.join(util.GOLD, c["dir"])
    base = [sys.executable, "-m", "sctagger_b200", "match_trie", "-lr", os.path.join(d, c["lr"]), "-sr",
            os.path.join(d, c["sr"]), "-mr", "2", "-t", "16"]
    r = subprocess.run(base, cwd=util.ROOT, capture_output=True)
    assert r.returncode == 0 and r.stdout == util.read_maybe_gz(os.path.join(d, c["out"]))
    r = subprocess.run(base + ["-bl", "15"], cwd=util.ROOT, capture_output=True)     # scTagger.py:760-761
    assert r.returncode != 0 and b"AssertionError" in r.stderr


def test_sctagger_py_entry_script_has_the_reference_command_line():
    """./scTagger.py is the reference's script name (scTagger.py:864-865): same sub-commands and flags."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "scTagger.py"), "match_trie", "-h"], capture_output=True, text=True)
    assert r.returncode == 0
    for flag in ("-lr", "-sr", "-mr", "-bl", "-m", "-t", "-p", "-o"):
        assert flag in r.stdout
    r = subprocess.run([sys.executable, os.path.join(root, "scTagger.py"), "match_trie", "-lr", "x"], capture_output=True, text=True)
    assert r.returncode == 2 and "-sr" in r.stderr          # argparse: a required flag is missing
