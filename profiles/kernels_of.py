"""Prints the per-kernel times of bench.py JSON lines: python profiles/kernels_of.py FILE..."""
import json
import sys

for path in sys.argv[1:]:
    line = [l for l in open(path) if l.strip().startswith("{")][-1]
    d = json.loads(line)
    ks = d["roofline"]["all_kernels"]
    print(path, "ms/step %.2f" % d["ms_per_step"], " ".join(
        "%s=%.2f" % (k.replace("plan_kernel<", "P<").replace("emit_objects", "E"), v["ms"]) for k, v in ks.items()))
