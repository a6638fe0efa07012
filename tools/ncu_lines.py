"""Per-source-line view of one kernel of an ncu report captured with --import-source on (kernels built with -lineinfo):
    python tools/ncu_lines.py gpurun_out/x.ncu-rep [kernel-regex] [top-N]
Aggregates `Instructions Executed` and warp-stall samples per CUDA source line (inlined code is attributed to its line)."""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
kern = sys.argv[2] if len(sys.argv) > 2 else "."
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name",
                      f"regex:{kern}", "--launch-count", "1"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = next(r for r in rows if r and r[0] == "Line No")
ie, smp = hdr.index("Instructions Executed"), hdr.index("# Samples")
seen, agg = set(), []
for r in rows:
    if r and r[0].isdigit() and (r[0], r[1]) not in seen:
        seen.add((r[0], r[1]))
        try:
            agg.append((int(r[0]), r[1].strip()[:100], int(r[ie]), int(r[smp])))
        except ValueError:
            pass
tot, ts = sum(a[2] for a in agg), sum(a[3] for a in agg)
print(f"# {rep}: {tot} warp instructions, {ts} samples, {len(agg)} source lines")
for a in sorted(agg, key=lambda a: -a[3])[:top]:
    print(f"{a[0]:5d} {100 * a[2] / tot:5.1f}% inst {100 * a[3] / ts:5.1f}% smp  {a[1]}")
