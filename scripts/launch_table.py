#!/usr/bin/env python
"""Groups an ncu launch list (--metrics gpu__time_duration.sum --csv) by kernel: launches, total / average time, share.

  python scripts/launch_table.py profiles/r02z_launches.csv > profiles/r02z_launches.txt
"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
h = rows[0]
ik, iv, iu, im = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit"), h.index("Metric Name")
agg = collections.OrderedDict()
for r in rows[1:]:
    if r[im] != "gpu__time_duration.sum":
        continue
    v = float(r[iv].replace(",", ""))
    us = v / 1000 if r[iu] in ("ns", "nsecond") else v if r[iu] in ("us", "usecond") else v * 1000
    a = agg.setdefault(re.sub(r"\(.*", "", r[ik]).strip(), [0, 0.0])
    a[0] += 1
    a[1] += us
tot = sum(a[1] for a in agg.values())
print("ncu --metrics gpu__time_duration.sum --clock-control none -k regex:compress, `python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu`")
print("(3 codecs x (3 warm-up + 2 timed) steps; cold-cache serialised times; the kernels' SHARE per codec is what agrees with the live CUDA-event split)")
print("launches     total us     avg us   share  kernel")
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{n:8d} {t:12.1f} {t / n:10.1f} {100 * t / tot:6.1f}%  {k}")
