"""Per-source-line instruction counts of decode_megakernel from an ncu report exported with
   ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > both.csv
Usage: python tools/mk_source_profile.py both.csv [steps] [layers]   (warp-instructions per warp per layer)"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
layers = int(sys.argv[3]) if len(sys.argv) > 3 else 22
sms, warps = 148, 12
cur, agg, hdr = None, [], None
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if len(r) > 8 and r[0] == "Line No":
        hdr = r
        continue
    if len(r) > 8 and r[0] not in ("", "Line No") and hdr:
        try:
            agg.append((cur, int(r[0]), r[1].strip()[:105], int(r[hdr.index("# Samples")]), int(r[hdr.index("Instructions Executed")])))
        except ValueError:
            pass
tot_i, tot_s = sum(a[4] for a in agg), sum(a[3] for a in agg) or 1
per = lambda i: i / steps / sms / warps / layers
print(f"total: {per(tot_i):.0f} warp-instructions per warp per layer; {tot_s} samples")
byfile = collections.Counter()
for a in agg:
    byfile[a[0]] += a[4]
for f, i in byfile.most_common(8):
    print(f"  {per(i):7.0f}  {f}")
print("top lines:")
for a in sorted(agg, key=lambda x: -x[4])[:60]:
    print(f"{per(a[4]):7.1f} inst {a[3] / tot_s * 100:5.1f}% smp  {a[0][:14]}:{a[1]:4d}  {a[2]}")
