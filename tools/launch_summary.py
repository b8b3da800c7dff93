#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel."""
import csv
import sys

path = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/launches.csv"
with open(path) as f:
    lines = [l for l in f if l.startswith('"')]
agg = {}
for row in csv.DictReader(lines):
    name = row["Kernel Name"].split("(")[0].replace("void ", "").replace("fm::<unnamed>::", "")[:48]
    v = float(row["Metric Value"].replace(",", ""))
    if row["Metric Unit"] in ("ns", "nsecond"):
        v /= 1e3
    a = agg.setdefault(name, [0, 0.0, row["Grid Size"], row["Block Size"]])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
print(f"{'kernel':50s} {'n':>4s} {'avg us':>10s} {'share':>7s}  grid block")
for k, (n, t, g, b) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{k:50s} {n:4d} {t / n:10.1f} {100 * t / tot:6.1f}%  {g} {b}")
