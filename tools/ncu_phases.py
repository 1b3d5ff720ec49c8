#!/usr/bin/env python3
"""Aggregate stall samples / executed instructions of an .ncu-rep (captured with --import-source on) over
source-line ranges of one file.  usage: python tools/ncu_phases.py report.ncu-rep file.cu name:lo-hi [name:lo-hi ...]"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
target = sys.argv[2]
ranges = []
for a in sys.argv[3:]:
    nm, r = a.split(":"); lo, hi = r.split("-"); ranges.append((nm, int(lo), int(hi)))
rows = list(csv.reader(out.splitlines()))
fname, hdr = None, None
tot_s = tot_i = 0
agg = {nm: [0, 0] for nm, _, _ in ranges}
other = {}
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No":
        hdr = r; continue
    if hdr and r[0].isdigit() and len(r) > 8 and r[2] == "-":
        d = dict(zip(hdr[4:], r[4:]))
        smp = int(d.get("# Samples", "0") or 0); ins = int(d.get("Instructions Executed", "0") or 0)
        tot_s += smp; tot_i += ins
        ln = int(r[0]); hit = False
        if fname == target:
            for nm, lo, hi in ranges:
                if lo <= ln <= hi:
                    agg[nm][0] += smp; agg[nm][1] += ins; hit = True; break
        if not hit:
            k = fname if fname != target else f"{fname}:other"
            o = other.setdefault(k, [0, 0]); o[0] += smp; o[1] += ins
print(f"total samples {tot_s}, warp instructions {tot_i}")
for nm, (s_, i_) in list(agg.items()) + sorted(other.items(), key=lambda kv: -kv[1][0]):
    print(f"{nm:28s} {100*s_/max(tot_s,1):5.1f}% samples  {100*i_/max(tot_i,1):5.1f}% instructions")
