#!/usr/bin/env python3
"""Per-source-line stall samples / instruction counts from an .ncu-rep captured with --import-source on.
usage: python tools/ncu_lines.py report.ncu-rep [top]"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(out.splitlines()))
fname, hdr, recs = None, None, []
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No":
        hdr = r; continue
    if hdr and r[0].isdigit() and len(r) > 8 and r[2] == "-":
        d = dict(zip(hdr[4:], r[4:]))
        recs.append((fname, int(r[0]), r[1].strip()[:90], int(d.get("# Samples", "0") or 0),
                     int(d.get("Instructions Executed", "0") or 0), d))
tot = sum(x[3] for x in recs) or 1
toti = sum(x[4] for x in recs) or 1
print(f"total samples {tot}, total warp instructions {toti}")
stalls = ["stall_barrier", "stall_long_sb", "stall_short_sb", "stall_math", "stall_wait", "stall_mio",
          "stall_not_selected", "stall_selected", "stall_branch_resolving", "stall_dispatch", "stall_lg", "stall_no_inst"]
agg = {s: sum(int(x[5].get(s, "0") or 0) for x in recs) for s in stalls}
print("stall totals:", {k: f"{100*v/tot:.1f}%" for k, v in agg.items() if v})
for f, ln, src, smp, ins, d in sorted(recs, key=lambda x: -x[3])[:top]:
    main = max(stalls, key=lambda s: int(d.get(s, "0") or 0))
    print(f"{f}:{ln:4d} {100*smp/tot:5.1f}% smp {100*ins/toti:5.1f}% inst  [{main}]  {src}")
