"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel.  usage: launch_summary.py launches.csv"""
import csv
import re
import sys
from collections import OrderedDict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = next(r for r in rows if "Kernel Name" in r)
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = OrderedDict()
for r in rows:
    if r is hdr or r[ki] == "Kernel Name":
        continue
    name = re.sub(r"\(.*", "", r[ki])[:60]
    v = float(r[vi].replace(",", ""))
    v = v / 1e6 if r[ui] in ("ns", "nsecond") else (v / 1e3 if r[ui] in ("us", "usecond") else v)
    c, t = agg.get(name, (0, 0.0))
    agg[name] = (c + 1, t + v)
total = sum(t for _, t in agg.values())
print("%-62s %6s %10s %7s" % ("kernel", "count", "total ms", "share"))
for name, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:20]:
    print("%-62s %6d %10.3f %6.1f%%" % (name, c, t, 100 * t / total))
