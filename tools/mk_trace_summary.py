import csv, sys, collections
rows = list(csv.DictReader(open(sys.argv[1])))
names = {0: "EMBED", 1: "QKV", 2: "ATTN", 3: "OPROJ", 4: "GATEUP", 5: "FC_GELU", 6: "DOWN", 7: "HEAD"}
mhz = float(sys.argv[2]) if len(sys.argv) > 2 else 1965.0
agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0])
prev_end = None
steps = max(int(r["step"]) for r in rows) + 1
for r in rows:
    if int(r["step"]) == 0:  # skip the cold first step
        prev_end = int(r["t_epilogue"]); continue
    t0, t1, t2, t3 = (int(r[k]) for k in ("t_start", "t_prologue", "t_main", "t_epilogue"))
    a = agg[names[int(r["kind"])]]
    a[0] += 1
    a[1] += (t0 - prev_end) / mhz   # barrier (end of previous phase -> start of this one)
    a[2] += (t1 - t0) / mhz         # prologue
    a[3] += (t2 - t1) / mhz         # main
    a[4] += (t3 - t2) / mhz         # epilogue
    if names[int(r["kind"])] == "ATTN" and int(r.get("n_stages", 0)) > 0:
        # attention stamps of CTA 0's unit (absolute clocks): q issued, K/V tile ready, scores done, softmax + P.V done
        q, kv, sc_, pv = (int(r[k]) for k in ("c_wait", "c_mma", "c_issue", "n_stages"))
        att = agg["ATTN:detail"]
        att[0] += 1; att[1] += (q - t0) / mhz; att[2] += (kv - q) / mhz; att[3] += (sc_ - kv) / mhz; att[4] += (pv - sc_) / mhz
        att[5] += (t3 - pv) / mhz
        prev_end = t3
        continue
    if 'c_wait' in r:
        a[5] += int(r['c_wait']) / mhz; a[6] += int(r['c_mma']) / mhz; a[7] += int(r['c_issue']) / mhz; a[8] += int(r['n_stages'])
    prev_end = t3
tot = 0
print(f"{'phase':8s} {'n/step':>6s} {'barrier':>8s} {'prologue':>9s} {'main':>8s} {'epilogue':>9s}   (us per phase instance; CTA 0)")
for k, a in agg.items():
    n = a[0]
    if k == "ATTN:detail":
        print(f"  attention unit of CTA 0: q issue {a[1]/n:.2f}  K/V wait {a[2]/n:.2f}  scores {a[3]/n:.2f}  softmax+PV {a[4]/n:.2f}  partial store {a[5]/n:.2f} us")
        continue
    extra = f"   warp0/stage: wait {a[5]/a[8]:.3f} mma {a[6]/a[8]:.3f} issue {a[7]/a[8]:.3f} us ({a[8]/n:.1f} stages)" if a[8] else ""
    print(f"{k:8s} {n/(steps-1):6.0f} {a[1]/n:8.2f} {a[2]/n:9.2f} {a[3]/n:8.2f} {a[4]/n:9.2f}{extra}")
    tot += sum(a[1:5])
print(f"per step: {tot/(steps-1):.1f} us")
