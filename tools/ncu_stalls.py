#!/usr/bin/env python
"""Stall reasons of one kernel by code region and the instructions that collect the most samples.
  python tools/ncu_stalls.py <report.ncu-rep> <lib.so> <kernel substring> [main file] [bucket] [top]"""
import collections, csv, io, os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from ncu_lines import sass_lines
rep, so, ksub = sys.argv[1:4]
main_file = sys.argv[4] if len(sys.argv) > 4 else "lmpc_warp_solve.cuh"
bucket = int(sys.argv[5]) if len(sys.argv) > 5 else 50
top = int(sys.argv[6]) if len(sys.argv) > 6 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]; ci = {n: i for i, n in enumerate(hdr)}
body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
sl = sass_lines(os.path.abspath(so), ksub)
n = min(len(sl), len(body))
reasons = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
reg = collections.defaultdict(collections.Counter)
inst = collections.Counter()
cur = 0; dm = {}
per = []
for k in range(n):
    loc, fn = sl[k][1], sl[k][3]
    if fn is None:
        if loc and loc[0] == main_file: cur = (loc[1] // bucket) * bucket
        key = f"body:{cur}"
    else:
        if fn not in dm: dm[fn] = "func " + subprocess.run(["c++filt", fn], capture_output=True, text=True).stdout.strip().split("(")[0][-28:]
        key = dm[fn]
    e = int(body[k][ci["Instructions Executed"]] or 0)
    inst[key] += e
    tot = 0
    for r in reasons:
        v = int(body[k][ci[r]] or 0); reg[key][r] += v; tot += v
    per.append((tot, k, key))
allr = collections.Counter()
for key in reg: allr.update(reg[key])
T = sum(allr.values())
print("all samples", T, {r[6:]: f"{100*v/T:.1f}%" for r, v in allr.most_common(9)})
print(f"{'region':40s} {'inst%':>6s} {'samp%':>6s}  top reasons (% of region samples)")
ti = sum(inst.values())
for key, c in sorted(reg.items(), key=lambda kv: -sum(kv[1].values()))[:28]:
    s = sum(c.values())
    print(f"{key:40s} {100*inst[key]/ti:6.2f} {100*s/T:6.2f}  " + ", ".join(f"{r[6:]}:{100*v/max(s,1):.0f}" for r, v in c.most_common(5)))
print("top instructions by samples:")
for tot, k, key in sorted(per, reverse=True)[:top]:
    rs = {r[6:]: int(body[k][ci[r]] or 0) for r in reasons}
    rs = ", ".join(f"{a}:{b}" for a, b in sorted(rs.items(), key=lambda kv: -kv[1])[:3] if b)
    loc = sl[k][1]
    print(f"  {tot:7d} {100*tot/T:5.2f}% {key:34s} {(loc[0]+':'+str(loc[1])) if loc else '?':28s} {sl[k][2][:60]:60s} {rs}")
