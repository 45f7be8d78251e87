#!/usr/bin/env python3
"""Per-kernel share of a step from an ncu launch list (ncu --metrics gpu__time_duration.sum --csv --log-file X).
   python tools/launch_shares.py launches.csv [first_launch last_launch]  -> JSON {kernel: {launches, us, share}}"""
import csv
import json
import re
import sys

rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith("==")]
rd = csv.DictReader(lines)
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "ns")
    us = v * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(unit, 1e-3)
    name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "").replace("hppb200::", "")
    rows.append((int(r["ID"]), name, us))
lo = int(sys.argv[2]) if len(sys.argv) > 2 else 0
hi = int(sys.argv[3]) if len(sys.argv) > 3 else 1 << 60
rows = [r for r in rows if lo <= r[0] <= hi]
tot = sum(r[2] for r in rows)
out = {}
for _, name, us in rows:
    d = out.setdefault(name, {"launches": 0, "us": 0.0})
    d["launches"] += 1
    d["us"] += us
for d in out.values():
    d["share"] = round(d["us"] / tot, 4)
    d["us"] = round(d["us"], 1)
print(json.dumps({"total_us": round(tot, 1), "launches": len(rows), "kernels": dict(sorted(out.items(), key=lambda kv: -kv[1]["us"]))}, indent=1))
