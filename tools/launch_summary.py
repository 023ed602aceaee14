#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per (kernel, grid): launches, mean / min us, share.
    python tools/launch_summary.py profiles/r01_v2_launches_cfg2.csv > profiles/r01_v2_launch_summary_cfg2.csv"""
import csv, re, sys
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
h = rows[0]
kn, gs, mv = h.index("Kernel Name"), h.index("Grid Size"), h.index("Metric Value")
agg = {}
for r in rows[1:]:
    name = re.sub(r"\(.*", "", r[kn]).replace("capdec::", "").replace("__nv_bfloat16", "bf16")
    agg.setdefault((name, r[gs]), []).append(float(r[mv]) / 1e3)
tot = sum(sum(v) for v in agg.values())
w = csv.writer(sys.stdout)
w.writerow(["kernel", "grid", "launches", "mean_us", "min_us", "share"])
for (name, grid), v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    w.writerow([name, grid, len(v), f"{sum(v) / len(v):.2f}", f"{min(v):.2f}", f"{sum(v) / tot:.3f}"])
