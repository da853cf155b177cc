#!/usr/bin/env python
"""Join a per-instruction dump (profiles/sass_hot.py) with the line table of the cubin it came
from (nvdisasm -g -gi) and sum samples / executed instructions / global loads per source line.

Usage: python profiles/sass_lines.py dump.csv file.cubin mangled-substring [top]"""
import collections
import csv
import re
import subprocess
import sys

dump, cubin, fn = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
rows = list(csv.reader(open(dump)))
hdr = rows[1]
ci = {c: i for i, c in enumerate(hdr)}
body = [r for r in rows[2:] if len(r) == len(hdr)]
n = len(body)
if n and body[: n // 2] and [r[ci["Source"]] for r in body[: n // 2]] == [r[ci["Source"]] for r in body[n // 2:]]:
    body = body[: n // 2]  # ncu prints the kernel twice when asked for one launch of several
txt = subprocess.run(["nvdisasm", "-g", "-gi", "-c", cubin], capture_output=True, text=True).stdout
import os
site = os.environ.get("SITE") == "1"
lines = []
cur = None
on = False
fresh = True
for ln in txt.split("\n"):
    if ln.startswith(".text."):
        on = fn in ln
        continue
    if not on:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        # an inline chain comes as several lines, innermost first: keep the innermost, or with
        # SITE=1 the innermost one outside the JSON toolkit (the planner's call site)
        loc = (m.group(1).split("/")[-1], int(m.group(2)))
        if fresh:
            cur = loc
            fresh = False
        elif site and cur[0] in ("nbrm_json.cuh", "nbrm_tape.cuh", "nbrm_plan.cuh", "nbrm_gen.cuh"):
            cur = loc
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
        lines.append(cur)
        fresh = True
print("instructions: dump %d, cubin %d" % (len(body), len(lines)))
agg = collections.defaultdict(lambda: [0.0, 0.0, 0.0, 0.0])


def f(x):
    try:
        return float(x)
    except ValueError:
        return 0.0


for k, r in enumerate(body):
    key = lines[k] if k < len(lines) and lines[k] else ("?", 0)
    a = agg[key]
    a[0] += f(r[ci["# Samples"]])
    a[1] += f(r[ci["Instructions Executed"]])
    a[2] += f(r[ci["L1 Tag Requests Global"]])
    a[3] += f(r[ci["stall_long_sb"]])
ts = sum(a[0] for a in agg.values()) or 1
ti = sum(a[1] for a in agg.values()) or 1
tl = sum(a[2] for a in agg.values()) or 1
print("%-26s %7s %7s %7s %7s" % ("line", "samp%", "inst%", "ld%", "longsb%"))
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%-26s %6.1f%% %6.1f%% %6.1f%% %6.1f%%" % ("%s:%d" % key, 100 * a[0] / ts, 100 * a[1] / ti, 100 * a[2] / tl,
                                                 100 * a[3] / ts))
