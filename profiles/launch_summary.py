"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv): python profiles/launch_summary.py launches.csv"""
import collections
import csv
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
d = collections.OrderedDict()
for row in csv.DictReader(lines):
    if row.get("Metric Name") == "gpu__time_duration.sum":
        us = float(row["Metric Value"]) / (1000.0 if row["Metric Unit"] in ("ns", "nsecond") else 1.0)
        d.setdefault(row["Kernel Name"].split("(")[0][-60:], []).append(us)
tot = sum(sum(v) for v in d.values())
for k, v in d.items():
    print(f"{k:60s} n={len(v):4d} mean={sum(v) / len(v):8.2f} us  min={min(v):7.2f} max={max(v):7.2f}  share={100 * sum(v) / tot:5.1f}%")
