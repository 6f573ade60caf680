#!/usr/bin/env python
"""SASS of the pod loop of a hot kernel: from the target of the last backward branch before the kernel's final barrier
to that branch.  usage: python tools/sass_excerpt.py <mangled name> <out file> [title]"""
import collections
import re
import subprocess
import sys

SO = "yoda_scheduler_b200/libyoda_b200.so"
name, out = sys.argv[1], sys.argv[2]
title = sys.argv[3] if len(sys.argv) > 3 else name
txt = subprocess.run(["cuobjdump", "-sass", "-fun", name, SO], capture_output=True, text=True).stdout
ins = []
for l in txt.splitlines():
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        ins.append((int(m.group(1), 16), m.group(2).strip()))
bars = [i for i, (_, t) in enumerate(ins) if t.startswith("BAR.SYNC")]
best = None
for i, (a, t) in enumerate(ins):
    m = re.search(r"BRA(?:\.U)?\s+(?:!?U?P\d,\s*)?(0x[0-9a-f]+)", t)
    if m and int(m.group(1), 16) < a and (not bars or i < bars[-1]):
        tgt = int(m.group(1), 16)
        if best is None or a - tgt > best[1] - best[0]:
            best = (tgt, a)
lo, hi = best
body = [(a, t) for a, t in ins if lo <= a <= hi]
cnt = collections.Counter(t.split()[1].split(".")[0] if t.startswith("@") else t.split()[0].split(".")[0] for _, t in body)
with open(out, "w") as f:
    f.write(f"# {title}\n# cuobjdump -sass -fun {name} {SO}\n# pod loop 0x{lo:x} .. 0x{hi:x}: {len(body)} instructions in the listing (all paths: hit, re-evaluation, exact rescale)\n")
    f.write("# mnemonics: " + ", ".join(f"{k} {v}" for k, v in cnt.most_common(14)) + "\n")
    for a, t in body:
        f.write(f"/*{a:04x}*/ {t} ;\n")
print(out, len(body), "instructions", dict(cnt.most_common(6)))
