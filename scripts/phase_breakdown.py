#!/usr/bin/env python
"""Instruction / stall-sample share per marked phase of a kernel source file.
usage: phase_breakdown.py report.ncu-rep kernel_regex source_file marker1 marker2 ...  (markers = substrings of source lines, in order)"""
import re, subprocess, sys, os
rep, kern, srcf = sys.argv[1:4]
marks = sys.argv[4:]
out = subprocess.run([sys.executable, os.path.join(os.path.dirname(__file__), "ncu_lines.py"), rep, kern, "2000"], capture_output=True, text=True).stdout
rows = []
for l in out.splitlines():
    m = re.match(r'\s*([\d.]+)%i\s+([\d.]+)%s\s+(\S+):\s*(\d+)\s(.*)', l)
    if m: rows.append((float(m[1]), float(m[2]), m[3], int(m[4]), m[5]))
print(out.splitlines()[0])
src = open(srcf).read().split('\n')
base = os.path.basename(srcf)
pts = [(mk[:28], [i + 1 for i, l in enumerate(src) if mk in l][0]) for mk in marks] + [("end", 10**9)]
for (n, a), (_, b) in zip(pts, pts[1:]):
    print(f"{n:28s} inst {sum(r[0] for r in rows if r[2]==base and a<=r[3]<b):5.1f}%  samples {sum(r[1] for r in rows if r[2]==base and a<=r[3]<b):5.1f}%")
others = {}
for r in rows:
    if r[2] != base:
        o = others.setdefault(r[2], [0, 0]); o[0] += r[0]; o[1] += r[1]
for k, v in others.items(): print(f"{k:28s} inst {v[0]:5.1f}%  samples {v[1]:5.1f}%")
print("top stall lines:")
for r in sorted(rows, key=lambda r: -r[1])[:12]: print("  ", r)
