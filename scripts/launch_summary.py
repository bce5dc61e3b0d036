#!/usr/bin/env python
"""Per-kernel totals of an ncu launch list (`--metrics gpu__time_duration.sum --csv`): count, total / mean time, share.
usage: launch_summary.py <launches.csv> [skip_first_n_launches]"""
import csv
import re
import sys
from collections import defaultdict

rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") == "gpu__time_duration.sum":
        rows.append((r["Kernel Name"], float(r["Metric Value"].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r["Metric Unit"], 1e-3), r["Grid Size"]))
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
rows = rows[skip:]
agg = defaultdict(lambda: [0, 0.0])
for name, us, grid in rows:
    short = re.sub(r"\(.*", "", name).replace("unnamed>::", "").replace("ace::", "")
    agg[short][0] += 1
    agg[short][1] += us
total = sum(v[1] for v in agg.values())
print(f"# {len(rows)} launches, {total / 1e3:.2f} ms in kernels (ncu per-launch times are cold-cache and serialised: shares, not absolutes)")
print("| kernel | launches | total ms | mean us | share |")
print("|---|---:|---:|---:|---:|")
for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"| {k} | {n} | {us / 1e3:.3f} | {us / n:.1f} | {100 * us / total:.1f} % |")
