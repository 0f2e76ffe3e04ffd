#!/usr/bin/env python
"""Per-stage instruction/sample shares of k_sweep from an .ncu-rep: python scripts/ncu_stage.py rep 'name:file:lo-hi,...'"""
import csv, subprocess, sys, io
rep = sys.argv[1]
spec = [s.split(":") for s in sys.argv[2].split(",")]
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-count", "1"], capture_output=True, text=True).stdout
cur, hd, agg = None, None, []
for r in csv.reader(io.StringIO(src)):
    if len(r) == 2 and r[0] == "File Path": cur = r[1].split("/")[-1]
    elif len(r) > 5 and r[0] == "Line No": hd = r
    elif hd and len(r) == len(hd) and r[0] != "": agg.append((cur, int(r[0]), int(r[6]), int(r[7]), int(r[8])))
ts, ti = sum(a[2] for a in agg) or 1, sum(a[3] for a in agg) or 1
used = set()
for name, f, rng in spec:
    lo, hi = [int(x) for x in rng.split("-")]
    sel = [a for a in agg if a[0].startswith(f) and lo <= a[1] <= hi]
    for a in sel: used.add((a[0], a[1]))
    s, i, t = sum(a[2] for a in sel), sum(a[3] for a in sel), sum(a[4] for a in sel)
    print(f"{name:10s} samples {100*s/ts:5.1f}%  warp-inst {100*i/ti:5.1f}% ({i:.3e})  thr/inst {t/max(1,i):5.1f}")
rest = [a for a in agg if (a[0], a[1]) not in used]
print(f"{'other':10s} samples {100*sum(a[2] for a in rest)/ts:5.1f}%  warp-inst {100*sum(a[3] for a in rest)/ti:5.1f}%")
for a in sorted(rest, key=lambda a: -a[2])[:8]: print("   ", a)
