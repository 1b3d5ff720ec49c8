#!/usr/bin/env python3
"""Histogram SASS opcodes inside the loops (backward branches) of one kernel.
usage: cuobjdump -sass lib.so | python tools/sass_loops.py <substring of mangled kernel name> [min_len]
Static counts only (both sides of forward branches are counted)."""
import re, sys, collections
pat = sys.argv[1]; min_len = int(sys.argv[2]) if len(sys.argv) > 2 else 200
ins = []; on = False
for line in sys.stdin:
    if "Function :" in line:
        on = pat in line
        continue
    if not on: continue
    m = re.match(r"\s+/\*([0-9a-f]+)\*/\s+(.*?);", line)
    if m:
        addr = int(m.group(1), 16); txt = m.group(2).strip()
        ins.append((addr, txt))
loops = []
for addr, txt in ins:
    m = re.search(r"BRA\S*\s+(?:\S+,\s*)?0x([0-9a-f]+)", txt)
    if m and "BRA" in txt:
        tgt = int(m.group(1), 16)
        if tgt < addr: loops.append((tgt, addr))
print("instructions:", len(ins), "loops:", [(hex(a), hex(b), (b - a) // 16) for a, b in loops if (b - a) // 16 >= min_len])
for a, b in loops:
    if (b - a) // 16 < min_len: continue
    c = collections.Counter()
    for addr, txt in ins:
        if a <= addr <= b:
            t = re.sub(r"^@!?U?P\w+\s+", "", txt)
            c[t.split()[0].split(".")[0]] += 1
    print(f"loop {hex(a)}..{hex(b)} len {(b-a)//16}:", dict(c.most_common()))
