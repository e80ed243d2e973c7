#!/usr/bin/env python
"""Aggregate an `ncu --page source --csv` dump by code region (relative address ranges).
usage: tools/ncu_regions.py src.csv start1:end1:name ...   (hex offsets relative to the kernel's first instruction)"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
base = int(data[0][0], 16)
regions = []
for a in sys.argv[2:]:
    s, e, n = a.split(":")
    regions.append((int(s, 16), int(e, 16), n))
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = {}
for r in data:
    off = int(r[0], 16) - base
    name = "other"
    for s, e, n in regions:
        if s <= off <= e:
            name = n
            break
    d = tot.setdefault(name, {"samples": 0, "inst": 0, **{k: 0 for k in stalls}})
    d["samples"] += int(r[ix["# Samples"]] or 0)
    d["inst"] += int(r[ix["Instructions Executed"]] or 0)
    for k in stalls:
        d[k] += int(r[ix[k]] or 0)
allsamp = sum(d["samples"] for d in tot.values())
for n, d in tot.items():
    top = sorted(((k, d[k]) for k in stalls), key=lambda kv: -kv[1])[:6]
    print(f"{n:10s} samples {d['samples']:8d} ({100*d['samples']/allsamp:5.1f}%) inst {d['inst']:12d}  samples/inst {d['samples']/max(d['inst'],1):.4f}  " +
          " ".join(f"{k[6:]}={v}" for k, v in top))
