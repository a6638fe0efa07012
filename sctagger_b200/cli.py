"""`scTagger.py match_trie` command line, flag-for-flag (reference scTagger.py:81-97, 140-143, 847-861).

    python -m sctagger_b200 match_trie -lr L.tsv.gz -sr S.tsv -mr 2 -bl 16 -t 16 -m 16 -o out.tsv.gz

Same flags, defaults and assertions as the reference.  Differences a user can see:
  * `-m` is honoured (it sizes the streamed chunks); the reference parses it and never reads it again.
  * `-p` is accepted and ignored — exactly what the reference's match_trie does (show_plot_match_trie is never called).
  * `--gpus N` / `--devices 0,1,..` (new, optional) shard the reads over several GPUs of this box.
  * log lines go to STDERR, so that without `-o` STDOUT holds the TSV rows only (the reference interleaves both).
Only the sub-command this package replaces is accepted; the other stages stay with the reference.
"""
from __future__ import annotations

import argparse
import contextlib
import sys


def build_parser() -> argparse.ArgumentParser:
    parser = argparse.ArgumentParser(prog="scTagger.py", description="scTagger pipeline! (match_trie on B200)")
    subparsers = parser.add_subparsers(dest="subcommand")
    subparsers.required = True
    p = subparsers.add_parser("match_trie")
    p.add_argument("-lr", "--long-read-segments", type=str, required=True, help="Long-read segments TSV file")
    p.add_argument("-sr", "--short-read-barcodes", type=str, required=True, help="Short-read barcode list TSV file")
    p.add_argument("-mr", "--max-error", default=2, type=int,
                   help="Maximum number of errors allowed for barcode matching. Default: 2")
    p.add_argument("-m", "--mem", default=16.0, type=float,
                   help="Maximum number of GB of RAM to be used. Default: 16.0")
    p.add_argument("-bl", "--barcode-length", default=16, type=int, help="Length of barcodes. Default: 16")
    p.add_argument("-t", "--threads", default=16, type=int,
                   help="Number of threads to use for searching. Default: 16")
    p.add_argument("-p", "--plotfile", default=None, type=str, help="Path of plot file. Default: no plotting")
    p.add_argument("-o", "--outfile", type=str, default=None,
                   help="Path to output file. Output file is gzipped. STDOUT is in normal text. Default: stdout")
    p.add_argument("--gpus", type=int, default=None, help="Number of GPUs to shard the reads over. Default: 1")
    p.add_argument("--devices", type=str, default=None, help="Comma separated CUDA device indices (overrides --gpus)")
    return parser


def parse_args(argv=None):
    args = build_parser().parse_args(argv)
    if args.subcommand == "match_trie":      # scTagger.py:140-143
        assert args.mem > 0
        assert args.barcode_length > 0
        assert args.barcode_length > args.max_error >= 0
        if args.devices is not None:
            args.devices = [int(x) for x in args.devices.split(",") if x != ""]
        elif args.gpus is not None:
            assert args.gpus >= 1
            args.devices = list(range(args.gpus))
    return args


def main(argv=None):
    from . import match_trie as mt
    args = parse_args(argv)
    # the reference prints the Namespace and its progress lines to STDOUT (scTagger.py:849, 604, 628, ...);
    # here they go to STDERR so that STDOUT output stays a clean TSV
    with contextlib.redirect_stdout(sys.stderr):
        print(args)
        if args.subcommand == "match_trie":
            data = mt.match_trie_bytes(args)
    mt.write_matches(args.outfile, data)
    return 0


if __name__ == "__main__":
    sys.exit(main())
