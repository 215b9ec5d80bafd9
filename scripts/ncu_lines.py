#!/usr/bin/env python
"""Per-source-line instruction counts / stall samples from an ncu report (cuda,sass view).
usage: ncu_lines.py report.ncu-rep kernel_regex [top_n]"""
import csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name",
                      "regex:" + kern], capture_output=True, text=True).stdout
fname, hdr, recs, on = None, None, [], True
for r in csv.reader(out.splitlines()):
    if len(r) >= 2 and r[0] == "File Path":
        fname = r[1].split("/")[-1]; hdr = None; continue
    if len(r) >= 2 and r[0] == "Function Name":
        on = kern.split("|")[0].replace("\\", "") in r[1] or True; continue
    if len(r) > 3 and r[0] == "Line No":
        hdr = r; ii = hdr.index("Instructions Executed"); si = hdr.index("# Samples"); continue
    if hdr and len(r) > max(ii, si) and r[0] and r[0].isdigit():
        try:
            ie = int(r[ii]); ss = int(r[si])
        except ValueError:
            continue
        if ie or ss:
            recs.append((ie, ss, fname, r[0], r[1].strip()[:110]))
ti = sum(r[0] for r in recs); ts = sum(r[1] for r in recs)
print(f"total inst {ti}  samples {ts}")
for ie, ss, f, ln, src in sorted(recs, key=lambda r: -r[0])[:top]:
    print(f"{100*ie/max(ti,1):5.1f}%i {100*ss/max(ts,1):5.1f}%s {f}:{ln:>4} {src}")
