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
# fine stamps (CTA 0, thread 0): where the time inside barrier / prologue / epilogue goes
fine = collections.defaultdict(lambda: [0] + [0.0] * 11)
prev_end = None
for r in rows:
    if "b_enter" not in r:
        break
    t0, t1, t2, t3 = (int(r[k]) for k in ("t_start", "t_prologue", "t_main", "t_epilogue"))
    if int(r["step"]) == 0 or prev_end is None:
        prev_end = t3
        continue
    g = {k: int(r[k]) for k in ("b_enter", "b_arrived", "b_done", "p_loaded", "p_reduced", "p_stored", "e_summed", "e_done")}
    a = fine[names[int(r["kind"])]]
    a[0] += 1
    if g["b_enter"]:
        a[1] += (g["b_enter"] - prev_end) / mhz      # first __syncthreads of the barrier (waiting for this CTA's other warps)
        a[2] += (g["b_arrived"] - g["b_enter"]) / mhz  # red.release (the CTA's stores must have been performed)
        a[3] += (g["b_done"] - g["b_arrived"]) / mhz   # poll until every CTA has arrived
        a[4] += (t0 - g["b_done"]) / mhz               # second __syncthreads + per-step set-up
    if g["p_loaded"]:
        a[5] += (g["p_loaded"] - t0) / mhz; a[6] += (g["p_reduced"] - g["p_loaded"]) / mhz; a[7] += (g["p_stored"] - g["p_reduced"]) / mhz
        a[8] += (t1 - g["p_stored"]) / mhz
    if g["e_summed"]:
        a[9] += (g["e_summed"] - t2) / mhz; a[10] += (g["e_done"] - g["e_summed"]) / mhz; a[11] += (t3 - g["e_done"]) / mhz
    prev_end = t3
print("fine (us): barrier = sync-in | red.release | poll | sync-out ; norm prologue = loads | reduce | normalise+store | sync ; "
      "epilogue = sum pieces | outputs of round 0 | rest")
for k, a in fine.items():
    n = a[0]
    print(f"{k:8s} {a[1]/n:6.2f} {a[2]/n:6.2f} {a[3]/n:6.2f} {a[4]/n:6.2f}   ; {a[5]/n:6.2f} {a[6]/n:6.2f} {a[7]/n:6.2f} {a[8]/n:6.2f}   ; {a[9]/n:6.2f} {a[10]/n:6.2f} {a[11]/n:6.2f}")
