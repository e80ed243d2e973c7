#!/usr/bin/env python
"""Aggregate one kernel of an `ncu --page source --csv` dump by code region (hex offsets from the kernel start).
usage: tools/ncu_regions.py src.csv kernel# start:end:name ..."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
kidx = int(sys.argv[2])
hdr_i = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
hdr = rows[hdr_i[kidx]]
ix = {h: i for i, h in enumerate(hdr)}
end = hdr_i[kidx + 1] - 1 if kidx + 1 < len(hdr_i) else len(rows)
data = [r for r in rows[hdr_i[kidx] + 1:end] if len(r) == len(hdr)]
base = int(data[0][0], 16)
regions = []
for a in sys.argv[3:]:
    s, e, n = a.split(":")
    regions.append((int(s, 16), int(e, 16), n))
stalls = [h for h in hdr if h.startswith("stall_") and "Not" not in h]
tot = {}
for r in data:
    off = int(r[0], 16) - base
    name = "other"
    for s, e, n in regions:
        if s <= off <= e:
            name = n
            break
    d = tot.setdefault(name, {"samples": 0, "inst": 0, "n": 0, **{k: 0 for k in stalls}})
    d["samples"] += int(r[ix["# Samples"]] or 0)
    d["inst"] += int(r[ix["Instructions Executed"]] or 0)
    d["n"] += 1
    for k in stalls:
        d[k] += int(r[ix[k]] or 0)
allsamp = sum(d["samples"] for d in tot.values())
allinst = sum(d["inst"] for d in tot.values())
for n, d in tot.items():
    top = sorted(((k[6:], d[k]) for k in stalls), key=lambda kv: -kv[1])[:4]
    print(f"{n:14s} static={d['n']:6d} exec={100*d['inst']/allinst:5.1f}% samples={100*d['samples']/allsamp:5.1f}% "
          f"ratio={(d['samples']/allsamp)/(d['inst']/allinst+1e-12):.2f} {top}")
