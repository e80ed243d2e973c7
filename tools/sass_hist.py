#!/usr/bin/env python
"""Opcode histogram per function of a `cuobjdump -sass` listing.  usage: tools/sass_hist.py file.sass [min]"""
import re, sys, collections
fn = None
h = collections.defaultdict(collections.Counter)
for line in open(sys.argv[1]):
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = re.sub(r"_ZN4ppcu\d+_GLOBAL__N__\w+?_cu_[0-9a-f]{8}", "", m.group(1))[:60]
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", line)
    if m and fn:
        h[fn][m.group(1).split(".")[0]] += 1
mn = int(sys.argv[2]) if len(sys.argv) > 2 else 20
for f, c in h.items():
    print(f, "total", sum(c.values()))
    print("   ", ", ".join(f"{k} {v}" for k, v in c.most_common() if v >= mn))
