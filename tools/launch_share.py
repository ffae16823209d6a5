#!/usr/bin/env python
"""ncu launch list (`--metrics gpu__time_duration.sum --csv --log-file X`) -> per-kernel share of the timed region.
usage: launch_share.py launches.csv [out.csv]"""
import collections
import csv
import re
import sys

rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith("==")]
rd = csv.DictReader(lines)
agg = collections.OrderedDict()
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\(.*", "", r["Kernel Name"])
    v = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "ns")
    us = v / 1e3 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1e3)
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += us
tot = sum(a[1] for a in agg.values())
n = sum(a[0] for a in agg.values())
out = [f"# total {tot / 1e3:.2f} ms over {n} launches (per-launch times are cold-cache and serialised: compare SHARES, not absolutes)",
       "kernel,launches,total_us,share,avg_us"]
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    out.append(f"\"{k}\",{c},{t:.1f},{t / tot:.4f},{t / c:.2f}")
txt = "\n".join(out)
print(txt)
if len(sys.argv) > 2:
    open(sys.argv[2], "w").write(txt + "\n")
