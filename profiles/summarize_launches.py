#!/usr/bin/env python
"""Per-kernel mean / share from an ncu launch list (--metrics gpu__time_duration.sum --csv)."""
import collections
import csv
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
agg = collections.OrderedDict()
for row in csv.DictReader(lines):
    v = float(row["Metric Value"].replace(",", ""))
    u = row["Metric Unit"]
    v = v if u == "ns" else v * 1e3 if u in ("us", "usecond") else v * 1e6 if u in ("ms", "msecond") else v
    agg.setdefault(row["Kernel Name"].split("(")[0][:48], []).append(v)
tot = sum(sum(v) for v in agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print("%-50s n=%3d mean=%9.1f us share=%5.1f%%" % (k, len(v), sum(v) / len(v) / 1e3, 100 * sum(v) / tot))
