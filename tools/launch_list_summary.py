#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name.
usage: python tools/launch_list_summary.py <launches.csv>"""
import csv
import re
import sys
from collections import defaultdict

rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 10]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot = defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    v = float(r[vi].replace(",", ""))
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3, "second": 1e6}.get(r[ui], 1.0)
    name = re.sub(r"\(.*", "", r[ki])[:70]
    tot[name][0] += 1
    tot[name][1] += v
total = sum(v[1] for v in tot.values())
print("total %.2f ms in %d launches" % (total / 1e3, sum(v[0] for v in tot.values())))
for name, (n, us) in sorted(tot.items(), key=lambda kv: -kv[1][1])[:30]:
    print("  %-70s %5d launches %12.1f us %5.1f%%" % (name, n, us, 100 * us / total))
