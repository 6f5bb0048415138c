"""Summarise an ncu launch list CSV (gpu__time_duration.sum) per kernel: count, mean us."""
import collections
import csv
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
agg = collections.OrderedDict()
for row in csv.DictReader(lines):
    if row.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(row["Metric Value"].replace(",", ""))
    u = row["Metric Unit"]
    v = v / 1e3 if u in ("nsecond", "ns") else v * 1e3 if u in ("msecond", "ms") else v
    a = agg.setdefault(row["Kernel Name"][:70], [0, 0.0])
    a[0] += 1
    a[1] += v
for k, (n, t) in agg.items():
    print(f"{n:4d} x {t / n:9.1f} us  {k}")
