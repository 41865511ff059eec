"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total and share."""
import csv
import sys
from collections import OrderedDict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 14 and r[0].isdigit()]
agg = OrderedDict()
for r in rows:
    name, grid, block, ns = r[4].split("(")[0], r[8], r[7], float(r[14])
    a = agg.setdefault(name, [0, 0.0, grid, block])
    a[0] += 1
    a[1] += ns
tot = sum(a[1] for a in agg.values())
print(f"# {sys.argv[1]}: {len(rows)} launches, {tot/1e6:.3f} ms total (cold-cache, serialised: compare shares)")
print(f"{'kernel':40s} {'launches':>8s} {'total_ms':>10s} {'avg_us':>10s} {'share':>7s}  last grid/block")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:40s} {a[0]:8d} {a[1]/1e6:10.3f} {a[1]/a[0]/1e3:10.1f} {a[1]/tot:7.1%}  {a[2]} {a[3]}")
