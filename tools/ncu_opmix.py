#!/usr/bin/env python3
"""Executed warp instructions and stall samples per SASS opcode from an .ncu-rep captured with --import-source on.
usage: python tools/ncu_opmix.py report.ncu-rep [iterations]   (iterations: divide the counts, e.g. Picard iterations)"""
import csv, subprocess, sys, collections
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
div = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
rows = list(csv.reader(out.splitlines()))
hdr = None
ins = collections.Counter(); smp = collections.Counter()
for r in rows:
    if not r:
        continue
    if "Source" in r and "# Samples" in r:
        hdr = r; continue
    if hdr and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        src = d["Source"].strip()
        toks = src.split()
        if not toks:
            continue
        op = toks[1] if toks[0].startswith("@") and len(toks) > 1 else toks[0]
        op = op.split(".")[0] + ("." + op.split(".")[1] if op.startswith(("MUFU", "SHFL", "BAR", "LDS", "STS")) and "." in op else "")
        try:
            ins[op] += int(d.get("Instructions Executed", "0") or 0); smp[op] += int(d.get("# Samples", "0") or 0)
        except ValueError:
            pass
ti = sum(ins.values()) or 1; ts = sum(smp.values()) or 1
print(f"total warp instructions {ti} ({ti/div:.1f} per unit), samples {ts}")
for op, c in ins.most_common(40):
    print(f"{op:14s} {c/div:12.1f} {100*c/ti:5.1f}% inst {100*smp[op]/ts:5.1f}% samples")
