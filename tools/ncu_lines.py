#!/usr/bin/env python
"""Per-source-line instruction counts and stall samples of the (single) kernel in an .ncu-rep captured with
--import-source on from a -lineinfo build.  usage: python tools/ncu_lines.py report.ncu-rep [top N]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
fname, hdr, lines = None, None, []
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        fname = r[1].split("/")[-1]
    elif len(r) > 8 and r[0] == "Line No":
        hdr = r
    elif hdr and len(r) == len(hdr) and r[2] == "-":   # a source line (its SASS rows follow with an address)
        try:
            lines.append((fname, int(r[0]), r[1].strip(), float(r[hdr.index("Instructions Executed")] or 0), float(r[hdr.index("# Samples")] or 0)))
        except ValueError:
            pass
tot_i = sum(l[3] for l in lines) or 1
tot_s = sum(l[4] for l in lines) or 1
print(f"total warp instructions {tot_i:.0f}, samples {tot_s:.0f}")
for f, n, src, ins, smp in sorted(lines, key=lambda l: -l[3])[:top]:
    print(f"{ins / tot_i * 100:5.1f}% inst {smp / tot_s * 100:5.1f}% smp  {f}:{n:<4d} {src[:130]}")
