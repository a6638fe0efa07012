"""`scTagger.py match_trie` command line, flag-for-flag (reference scTagger.py:81-97, 140-143, 847-861).

    python -m sctagger_b200 match_trie -lr L.tsv.gz -sr S.tsv -mr 2 -bl 16 -t 16 -m 16 -o out.tsv.gz

Same flags, defaults and assertions as the reference.  Differences a user can see:
  * `-m` is honoured (it sizes the streamed chunks); the reference parses it and never reads it again.
  * `-p` is accepted and ignored — exactly what the reference's match_trie does (show_plot_match_trie is never called).
  * `--gpus N` / `--devices 0,1,..` (new, optional) shard the reads over several GPUs of this box.
  * log lines go to STDERR, so that without `-o` STDOUT holds the TSV rows only (the reference interleaves both).
`extract_lr_bc` (reference scTagger.py:25-49, 100-128) is accepted with its own flags as well: its adapter alignment runs
on the GPU (`-p` plotting is not part of this path and is ignored).  The other stages stay with the reference.
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
    q = subparsers.add_parser("extract_lr_bc")          # scTagger.py:25-49
    q.add_argument("-r", "--reads", nargs="+", type=str, required=True, help="Space separated paths to reads in FASTQ")
    q.add_argument("-g", "--ranges", nargs="+", type=str, default=list(),
                   help="Space separated ranges of where the SR adapter should be found on the LRs, e.g. f20:40 r1:30 "
                        "(1-indexed, inclusive; f = from the start on the forward strand, r = from the end on the "
                        "reverse strand). Default: Detect from data.")
    q.add_argument("-z", "--gzipped", dest="gzipped", action="store_true",
                   help="Indicate input is gzipped. Default: Assume input is gzipped if it ends with \".gz\".")
    q.add_argument("-t", "--threads", default=1, type=int, help="Number of threads. Default: 1")
    q.add_argument("-sa", "--short-read-adapter", type=str, default="CTACACGACGCTCTTCCGATCT",
                   help="Short-read adapter. Default: CTACACGACGCTCTTCCGATCT")
    q.add_argument("-o", "--outfile", type=str, default=None,
                   help="Path to output file. Output file is gzipped. STDOUT is in normal text. Default: stdout")
    q.add_argument("-p", "--plotfile", type=str, default=None, help="Path to plot file. Default: no plotting")
    q.add_argument("--num-bp-after", type=int, default=20,
                   help="Number of bases after the end of the SR adapter alignment to generate. Default: 20")
    q.add_argument("--gpus", type=int, default=None, help="Number of GPUs to shard the reads over. Default: 1")
    q.add_argument("--devices", type=str, default=None, help="Comma separated CUDA device indices (overrides --gpus)")
    w = subparsers.add_parser("extract_sr_bc_from_lr")      # scTagger.py:67-79
    w.add_argument("-i", "--input", type=str, required=True, help="Input TSV file generated from extract_lr_bc step")
    w.add_argument("-o", "--outfile", type=str, default=None, help="Path to output file. Default: STDOUT")
    w.add_argument("-wl", "--barcode-whitelist", type=str, required=True,
                   help="Path to TXT barcode whiltelist such as 10x Chromimum v3 3M-february-2018.txt.gz barcode file")
    w.add_argument("--thresh", type=float, default=0.005,
                   help="Percentage theshold required per step to continue adding read barcodes. Default: 0.005")
    w.add_argument("--step-size", type=int, default=1000,
                   help="Number of barcodes processed at a time and whose sum is used to check against the theshold. "
                        "Default: 1000")
    w.add_argument("--max-barcode-cnt", type=int, default=25_000, help="Max number of barcodes to keep. Default: 25000")
    w.add_argument("--gpus", type=int, default=None, help="(one GPU is used; the table is not sharded)")
    w.add_argument("--devices", type=str, default=None, help="CUDA device index")
    return parser


def _devices(args):
    if args.devices is not None:
        args.devices = [int(x) for x in args.devices.split(",") if x != ""]
    elif args.gpus is not None:
        assert args.gpus >= 1
        args.devices = list(range(args.gpus))


def parse_args(argv=None):
    args = build_parser().parse_args(argv)
    if args.subcommand == "extract_lr_bc":   # scTagger.py:100-128
        from .extract_lr_bc import parse_ranges_flag
        assert 0 < args.num_bp_after
        args.ranges = parse_ranges_flag(args.ranges)
        assert args.threads > 0
        _devices(args)
    if args.subcommand == "extract_sr_bc_from_lr":   # scTagger.py:135-138
        assert 0 <= args.thresh <= 1
        assert 0 < args.step_size
        assert 0 < args.max_barcode_cnt
        _devices(args)
    if args.subcommand == "match_trie":      # scTagger.py:140-143
        assert args.mem > 0
        assert args.barcode_length > 0
        assert args.barcode_length > args.max_error >= 0
        _devices(args)
    return args


def main(argv=None):
    from . import match_trie as mt
    args = parse_args(argv)
    if args.subcommand == "extract_lr_bc":
        # the reference logs this stage to STDERR already (scTagger.py:163, 213, 249, ...); only the Namespace goes to STDOUT
        from . import extract_lr_bc as lr
        print(args, file=sys.stderr)
        lr.extract_lr_bc(args)
        return 0
    if args.subcommand == "extract_sr_bc_from_lr":
        from . import extract_sr_bc_from_lr as wl
        with contextlib.redirect_stdout(sys.stderr):
            print(args)
            data = wl.extract_sr_bc_from_lr_bytes(args)
        if args.outfile:
            import gzip
            with gzip.open(args.outfile, "wb", compresslevel=mt.GZIP_LEVEL) as f:
                f.write(data)
        else:
            sys.stdout.buffer.write(data)
            sys.stdout.buffer.flush()
        return 0
    # the reference prints the Namespace and its progress lines to STDOUT (scTagger.py:849, 604, 628, ...);
    # here they go to STDERR so that STDOUT output stays a clean TSV
    if args.outfile:
        with contextlib.redirect_stdout(sys.stderr):
            print(args)
            mt.match_trie(args)
        return 0
    with contextlib.redirect_stdout(sys.stderr):
        print(args)
        data = mt.match_trie_bytes(args)
    mt.write_matches(None, data)
    return 0


if __name__ == "__main__":
    sys.exit(main())
