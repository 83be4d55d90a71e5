#!/usr/bin/env python
"""ncu_launch_summary.py -- per-kernel totals of an ncu launch list (`--metrics gpu__time_duration.sum --csv`):
launch count, mean / min / max duration and the kernel's SHARE of the summed GPU time (per-launch times under ncu are
cold-cache and serialised: the share is what must agree with bench.py's event timing, not the absolute).
usage: scripts/ncu_launch_summary.py launches.csv > summary.json"""
import csv
import json
import re
import sys

rows = []
with open(sys.argv[1], newline="") as f:
    lines = [l for l in f if not l.startswith("==")]
rd = csv.DictReader(lines)
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "ns")
    ms = v * {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "nsecond": 1e-6, "ms": 1.0, "msecond": 1.0, "s": 1e3, "second": 1e3}.get(unit, 1e-6)
    name = re.sub(r"\(.*", "", r["Kernel Name"]).strip()
    rows.append((name, ms))
tot = sum(ms for _, ms in rows) or 1.0
agg = {}
for name, ms in rows:
    a = agg.setdefault(name, [])
    a.append(ms)
out = {"launches": len(rows), "total_ms": tot, "kernels": []}
for name, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    out["kernels"].append({"kernel": name, "launches": len(v), "mean_ms": sum(v) / len(v), "min_ms": min(v), "max_ms": max(v), "share": sum(v) / tot})
print(json.dumps(out, indent=1))
