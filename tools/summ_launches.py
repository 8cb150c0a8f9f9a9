#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel."""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = [i for i, r in enumerate(rows) if r[0] == "ID"][0]
H = rows[hdr]
ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
d = collections.OrderedDict()
for r in rows[hdr + 1:]:
    n = r[ki].split("(")[0][:60]
    v = float(r[vi].replace(",", ""))
    v = v / 1e3 if r[ui] == "ns" else (v * 1e3 if r[ui] == "ms" else v)
    d.setdefault(n, []).append(v)
for n, v in d.items():
    print("%-62s n=%3d avg=%9.1f us max=%9.1f" % (n, len(v), sum(v) / len(v), max(v)))
