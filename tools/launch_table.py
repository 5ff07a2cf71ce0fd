#!/usr/bin/env python
"""Aggregate a `bench.py --dump-launches` JSON by (kernel, shape tag): launches/step, ms/step, TFLOP/s or GB/s.

    python tools/launch_table.py gpurun_out/launches.json [steps]
"""
import json
import sys
from collections import OrderedDict

rows = json.load(open(sys.argv[1]))
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
agg = OrderedDict()
for r in rows:
    k = (r["name"].replace("seva_", ""), r["tag"])
    a = agg.setdefault(k, dict(ms=0.0, n=0, fl=0.0, by=0.0))
    a["ms"] += r["ms"]
    a["n"] += 1
    a["fl"] += r["flops"]
    a["by"] += r["bytes"]
tot = sum(a["ms"] for a in agg.values())
print(f"total {tot / steps:.2f} ms/step over {len(rows) // steps} launches/step")
print(f"{'kernel':18s} {'shape':40s} {'n/step':>6s} {'ms/step':>8s} {'share':>6s} {'us/launch':>9s} {'TF/s':>7s} {'GB/s':>7s}")
for (name, tag), a in sorted(agg.items(), key=lambda kv: -kv[1]["ms"]):
    tf = a["fl"] / a["ms"] / 1e9 if a["fl"] else 0
    gb = a["by"] / a["ms"] / 1e6 if a["by"] else 0
    print(f"{name:18s} {tag:40s} {a['n'] / steps:6.1f} {a['ms'] / steps:8.3f} {a['ms'] / tot:6.3f} "
          f"{a['ms'] / a['n'] * 1e3:9.1f} {tf:7.1f} {gb:7.1f}")
