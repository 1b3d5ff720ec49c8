#!/usr/bin/env python3
"""SASS instruction count per source line of one kernel (needs -lineinfo).
usage: python tools/sass_lines.py obj.o <substring of mangled kernel name> [top]"""
import collections, os, re, subprocess, sys, tempfile
obj, pat = os.path.abspath(sys.argv[1]), sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
with tempfile.TemporaryDirectory() as d:
    subprocess.run(["cuobjdump", "-xelf", "all", obj], cwd=d, capture_output=True)
    cub = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
    out = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, cub)], capture_output=True, text=True).stdout
cnt, ops, cur, on = collections.Counter(), collections.Counter(), None, False
for l in out.splitlines():
    m = re.match(r"\s*\.section\s+\.text\.(\S+?),", l)
    if m:
        on = pat in m.group(1); continue
    if not on:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m = re.match(r"\s+/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", l)
    if m:
        cnt[cur] += 1; ops[m.group(1).split(".")[0]] += 1
tot = sum(cnt.values())
print("total SASS instructions:", tot)
print("opcodes:", ", ".join(f"{k} {v}" for k, v in ops.most_common(18)))
for k, v in sorted(cnt.items(), key=lambda x: -x[1])[:top]:
    print(f"{k[0] if k else '?'}:{k[1] if k else 0:4d}  {v}")
