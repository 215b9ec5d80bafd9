#!/usr/bin/env python
"""Per-source-line instruction counts and stall samples of one kernel in an ncu report.

ncu's csv source page is per SASS instruction and carries no line numbers, so the SASS rows are joined, in order,
with `nvdisasm -g` of the same kernel taken from the library the report was made with (build with -lineinfo).

  ncu_lines.py report.ncu-rep <kernel-regex> [lib.so] [top N]
"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

rep, pat = sys.argv[1], sys.argv[2]
lib = sys.argv[3] if len(sys.argv) > 3 else os.path.join(os.path.dirname(__file__), "..", "hipcomp-core_b200", "lib", "libhipcomp.so")
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40

out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{pat}"],
                     capture_output=True, text=True).stdout
blocks = out.split('"Kernel Name",')
# ncu matches the bare function name only: NCU_LINES_MATCH=<substring> picks among template instantiations
want = os.environ.get("NCU_LINES_MATCH")
blk = next((b for b in blocks[1:] if want is None or want in b.splitlines()[0]), None)
if blk is None:
    sys.exit("no kernel in the report matches")
kname = blk.splitlines()[0]
rows = list(csv.reader(io.StringIO("\n".join(blk.splitlines()[1:]))))
h = rows[0]
iS, iI, iN = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
stall_cols = [(i, n) for i, n in enumerate(h) if n.startswith("stall_") and "Not Issued" not in n]
sass = [r for r in rows[1:] if len(r) == len(h)]

tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
func = re.search(r"(?:void )?(?:\w+::)*(\w+)\s*[<(]", kname.strip(chr(34))).group(1)
lines = None
for f in os.listdir(tmp):
    dis = subprocess.run(["nvdisasm", "-gi", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
    # split by .text sections; pick the one whose demangled template args match
    secs = re.split(r"^\s*\.text\.", dis, flags=re.M)
    for s in secs[1:]:
        name = s.split(":", 1)[0]
        if func not in name:
            continue
        cur, seq, fresh = None, [], True
        for ln in s.splitlines():
            m = re.search(r'//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', ln)
            if m:
                # -gi prints the inline chain innermost first; keep the outermost frame (the kernel's own line),
                # or with NCU_LINES_INNER=1 the innermost one (the helper's line)
                if os.environ.get("NCU_LINES_INNER"):
                    if fresh:
                        cur = (os.path.basename(m.group(1)), int(m.group(2)))
                    fresh = False
                elif m.group(3):
                    cur = (os.path.basename(m.group(3)), int(m.group(4)))
                else:
                    cur = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
                seq.append(cur)
                fresh = True
        if len(seq) == len(sass):
            lines = seq
            break
    if lines:
        break
if lines is None:
    sys.exit(f"no disassembled function named {func} with {len(sass)} instructions")

agg = {}
tot_i = tot_s = 0
for r, loc in zip(sass, lines):
    a = agg.setdefault(loc, [0, 0, {}])
    a[0] += int(r[iI])
    a[1] += int(r[iN])
    tot_i += int(r[iI])
    tot_s += int(r[iN])
    for i, n in stall_cols:
        v = int(r[i] or 0)
        if v:
            a[2][n] = a[2].get(n, 0) + v
print(f"{kname[:100]}\ninstructions {tot_i}, samples {tot_s}")
src_cache = {}
for loc, (ni, ns, st) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    txt = ""
    if loc:
        p = os.path.join(os.path.dirname(__file__), "..", "hipcomp-core_b200", "csrc", loc[0])
        if os.path.exists(p):
            src_cache.setdefault(p, open(p).read().splitlines())
            txt = src_cache[p][loc[1] - 1].strip()[:90]
    tops = ", ".join(f"{k[6:]} {v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
    print(f"{100 * ni / tot_i:5.1f}% inst {100 * ns / max(tot_s, 1):5.1f}% smpl  {loc}  {txt}   [{tops}]")
