#!/usr/bin/env python
"""Who calls one hot inlined line: samples and global loads of the instructions whose innermost
frame is FILE:LINE, grouped by the next frame out.
Usage: python profiles/sass_callers.py dump.csv file.cubin mangled-substring FILE LINE [top]"""
import collections, csv, re, subprocess, sys
dump, cubin, fn, hf, hl = sys.argv[1:6]
top = int(sys.argv[6]) if len(sys.argv) > 6 else 20
rows = list(csv.reader(open(dump)))
hdr = rows[1]
ci = {c: i for i, c in enumerate(hdr)}
body = [r for r in rows[2:] if len(r) == len(hdr)]
txt = subprocess.run(["nvdisasm", "-g", "-gi", "-c", cubin], capture_output=True, text=True).stdout
chains, cur, on = [], [], False
for ln in txt.split("\n"):
    if ln.startswith(".text."):
        on = fn in ln
        continue
    if not on:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur.append((m.group(1).split("/")[-1], int(m.group(2))))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
        chains.append(cur)
        cur = []
last = []
for k, c in enumerate(chains):
    if c:
        last = c
    else:
        chains[k] = last


def f(x):
    try:
        return float(x)
    except ValueError:
        return 0.0


agg = collections.defaultdict(lambda: [0.0, 0.0])
for k, r in enumerate(body[:len(chains)]):
    c = chains[k]
    if not c or c[0] != (hf, int(hl)):
        continue
    key = c[1] if len(c) > 1 else ("?", 0)
    agg[key][0] += f(r[ci["# Samples"]])
    agg[key][1] += f(r[ci["L1 Tag Requests Global"]])
ts = sum(f(r[ci["# Samples"]]) for r in body) or 1
tl = sum(f(r[ci["L1 Tag Requests Global"]]) for r in body) or 1
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%-24s samp %5.1f%%  ld %5.1f%%" % ("%s:%d" % key, 100 * a[0] / ts, 100 * a[1] / tl))
