#!/usr/bin/env python
"""List the loops (backward branches) of one kernel in a built library with their SASS instruction mix.
usage: tools/sass_loops.py <lib.so> <kernel-name-substring>"""
import re, subprocess, sys
lib, pat = sys.argv[1], sys.argv[2]
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
ins, on = [], False
for l in out.split("\n"):
    if "Function :" in l:
        on = pat in l
        continue
    if not on:
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        ins.append((int(m.group(1), 16), m.group(2)))
print("instructions:", len(ins))
def mix(body):
    ops = {}
    for _, t in body:
        f = t.split()
        op = (f[1] if f[0].startswith("@") else f[0]).split(".")[0]
        ops[op] = ops.get(op, 0) + 1
    return sorted(ops.items(), key=lambda kv: -kv[1])[:9]
print("all:", mix(ins))
for a, t in ins:
    m = re.search(r"BRA\S*\s+.*?(0x[0-9a-f]+)", t)
    if m and int(m.group(1), 16) < a:
        tgt = int(m.group(1), 16)
        body = [x for x in ins if tgt <= x[0] <= a]
        print(hex(tgt), "..", hex(a), len(body), mix(body))
