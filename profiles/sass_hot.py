#!/usr/bin/env python
"""Cut the `ncu --page source --csv` dump of one launch down to the columns the hot-spot
notes use (run on the GPU box, where the report is).

Usage: ncu -i x.ncu-rep --page source --csv --launch-skip N --launch-count 1 | python profiles/sass_hot.py > out.csv"""
import csv
import sys

KEEP = ["Address", "Source", "# Samples", "Instructions Executed", "Avg. Threads Executed", "stall_long_sb",
        "stall_short_sb", "stall_wait", "stall_no_inst", "stall_branch_resolving", "L1 Tag Requests Global"]
rows = csv.reader(sys.stdin)
w = csv.writer(sys.stdout)
hdr = None
for r in rows:
    if hdr is None:
        if "Source" in r and "Address" in r:
            hdr = [r.index(k) if k in r else -1 for k in KEEP]
            w.writerow(KEEP)
        else:
            w.writerow(r[:2])
        continue
    if len(r) < 3:
        hdr = None  # next kernel's preamble
        w.writerow(r[:2])
        continue
    w.writerow([r[i] if 0 <= i < len(r) else "" for i in hdr])
