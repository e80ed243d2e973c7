#!/usr/bin/env python
"""Split one kernel of an `ncu --page source --csv` dump into runs of instructions with the same execution count and
print their share of executed instructions and of stall samples.  usage: tools/ncu_segments.py src.csv [kernel#] [min%]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
kidx = int(sys.argv[2]) if len(sys.argv) > 2 else 0
minpct = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
hdr_i = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
hdr = rows[hdr_i[kidx]]
ix = {h: i for i, h in enumerate(hdr)}
end = hdr_i[kidx + 1] - 1 if kidx + 1 < len(hdr_i) else len(rows)
data = [r for r in rows[hdr_i[kidx] + 1:end] if len(r) == len(hdr)]
print(rows[hdr_i[kidx] - 1][1][:100])
base = int(data[0][0], 16)
stalls = [h for h in hdr if h.startswith('stall_') and 'Not' not in h]
segs, cur = [], None
for r in data:
    e = int(r[ix['Instructions Executed']] or 0)
    off = int(r[0], 16) - base
    if cur is None or abs(e - cur['e']) > 0.02 * max(e, cur['e'], 1):
        cur = {'start': off, 'e': e, 'n': 0, 'samp': 0, 'exec': 0, 'st': {}}
        segs.append(cur)
    cur['n'] += 1
    cur['samp'] += int(r[ix['# Samples']] or 0)
    cur['exec'] += e
    cur['end'] = off
    for k in stalls:
        cur['st'][k] = cur['st'].get(k, 0) + int(r[ix[k]] or 0)
tot = sum(s['samp'] for s in segs)
te = sum(s['exec'] for s in segs)
print('instructions', len(data), 'executed', te, 'samples', tot)
for s in segs:
    if s['samp'] > minpct / 100 * tot:
        top = sorted(s['st'].items(), key=lambda kv: -kv[1])[:4]
        print(f"{s['start']:6x}-{s['end']:6x} n={s['n']:5d} e={s['e']:8d} exec={100*s['exec']/te:5.1f}% samp={100*s['samp']/tot:5.1f}%",
              [(k[6:], v) for k, v in top])
