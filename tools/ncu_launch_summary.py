#!/usr/bin/env python
"""Summarise an ncu launch list (`--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv`)
per kernel: launches, total device time, share, DRAM bytes per launch.  Prints CSV.

    python tools/ncu_launch_summary.py gpurun_out/prof_r1b/launches.csv [skip_launches] > profiles/r1b_ncu_launch_summary.csv
"""
import csv
import re
import sys
from collections import OrderedDict

path = sys.argv[1]
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
rows = {}
with open(path, newline="") as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    d = rows.setdefault(int(r["ID"]), {"name": r["Kernel Name"]})
    v = float(r["Metric Value"].replace(",", ""))
    u = r["Metric Unit"]
    if r["Metric Name"] == "gpu__time_duration.sum":
        d["us"] = v / 1e3 if u in ("ns", "nsecond") else (v if u in ("us", "usecond") else v * 1e3)
    else:
        scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
        d["rd" if "read" in r["Metric Name"] else "wr"] = v * scale


def short(n):
    n2 = re.sub(r"void |seva::|\(anonymous namespace\)::|<unnamed>::", "", n)
    return n2.split("(")[0].split("<")[0][:60]


agg = OrderedDict()
for i in sorted(rows):
    if i < skip:
        continue
    d = rows[i]
    a = agg.setdefault(short(d["name"]), dict(n=0, us=0.0, rd=0.0, wr=0.0))
    a["n"] += 1
    a["us"] += d.get("us", 0.0)
    a["rd"] += d.get("rd", 0.0)
    a["wr"] += d.get("wr", 0.0)
tot = sum(a["us"] for a in agg.values())
w = csv.writer(sys.stdout)
w.writerow(["kernel", "launches", "total_ms", "share", "avg_us", "dram_read_MB_per_launch", "dram_write_MB_per_launch"])
for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["us"]):
    w.writerow([k, a["n"], f"{a['us'] / 1e3:.3f}", f"{a['us'] / tot:.4f}", f"{a['us'] / a['n']:.1f}",
                f"{a['rd'] / a['n'] / 1e6:.2f}", f"{a['wr'] / a['n'] / 1e6:.2f}"])
