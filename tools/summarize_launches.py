"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import csv
import re
import sys
from collections import defaultdict

rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith("==")]
rd = csv.DictReader(lines)
tot = defaultdict(lambda: [0, 0.0])
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = r["Kernel Name"]
    name = re.sub(r"\(.*", "", name)
    val = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    ms = val / 1e6 if unit in ("ns", "nsecond") else (val / 1e3 if unit in ("us", "usecond") else val)
    tot[name][0] += 1
    tot[name][1] += ms
total = sum(v[1] for v in tot.values())
print("%-70s %6s %10s %7s" % ("kernel", "count", "ms", "share"))
for k, v in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print("%-70s %6d %10.3f %6.1f%%" % (k[:70], v[0], v[1], 100 * v[1] / total))
print("%-70s %6d %10.3f" % ("TOTAL", sum(v[0] for v in tot.values()), total))
