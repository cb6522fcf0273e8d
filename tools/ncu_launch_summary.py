"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel name.  Usage: python tools/ncu_launch_summary.py file.csv [name-width]"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
width = int(sys.argv[2]) if len(sys.argv) > 2 else 90
hdr = [i for i, r in enumerate(rows) if r[0] == "ID"][0]
H = rows[hdr]
ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[hdr + 1:]:
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    v = v / 1000 if r[ui] == "ns" else (v * 1000 if r[ui] == "ms" else v)
    agg[r[ki][:width]][0] += 1
    agg[r[ki][:width]][1] += v
tot = sum(a[1] for a in agg.values())
for n, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{a[1] / tot * 100:5.1f}% {a[0]:6d} launches {a[1] / a[0]:9.1f} us each  {n}")
print(f"total {tot:.1f} us")
