#!/usr/bin/env python
"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list.
Usage: python profiles/summarize_launches.py profiles/r01_launches_v5.csv"""
import collections
import csv
import re
import sys

rows = []
with open(sys.argv[1], newline="") as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = r["Kernel Name"]
    m = re.search(r"plan_kernel<[^>]*?(Gen\w+)>", name)
    name = "plan_kernel<%s>" % m.group(1) if m else name.split("(")[0].split("::")[-1]
    v = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1.0)
    rows.append((name, v))
tot = collections.defaultdict(float)
cnt = collections.Counter()
for n, v in rows:
    tot[n] += v
    cnt[n] += 1
total = sum(tot.values())
for n in sorted(tot, key=tot.get, reverse=True):
    print("%-44s launches %3d  total %9.1f us  avg %8.1f us  share %5.1f%%"
          % (n, cnt[n], tot[n], tot[n] / cnt[n], 100 * tot[n] / total))
