#!/usr/bin/env python
"""Attribute the stall samples and executed instructions of one kernel (ncu --page source --csv) to source lines, using
the line table of the same build (nvdisasm -g -c of the object's cubin).
usage: tools/ncu_lines.py src.csv kernel-substring nvdisasm.sass [top]"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
sub, sass, top = sys.argv[2], sys.argv[3], int(sys.argv[4]) if len(sys.argv) > 4 else 40
# line table: offset -> (file, line, inlined-at chain text)
# (nvdisasm -gi lists the inline chain of an instruction innermost first; the key is its outermost DEPTH frames)
DEPTH = int(sys.argv[5]) if len(sys.argv) > 5 else 2
chain, fn, table, fresh = [], None, {}, True
for l in open(sass):
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        if fresh:
            chain, fresh = [], False
        chain.append(m.group(1).split('/')[-1].replace('rs255_solve_core.h', 'core') + ':' + m.group(2))
        continue
    m = re.match(r'\s*\.text\.(\S+):', l)
    if m:
        fn = m.group(1)
        continue
    m = re.match(r'\s+/\*([0-9a-f]+)\*/\s+(\S.*?);', l)
    if m:
        fresh = True
        if fn and sub in fn:
            table[int(m.group(1), 16)] = (' < '.join(chain[-DEPTH:]), m.group(2))
hdr_i = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
for k, hi in enumerate(hdr_i):
    name = rows[hi - 1][0] if hi else ''
    hdr = rows[hi]
    ix = {h: i for i, h in enumerate(hdr)}
    end = hdr_i[k + 1] - 1 if k + 1 < len(hdr_i) else len(rows)
    data = [r for r in rows[hi + 1:end] if len(r) == len(hdr)]
    if not data:
        continue
    base = int(data[0][0], 16)
    ex, sm = collections.Counter(), collections.Counter()
    stall = collections.defaultdict(collections.Counter)
    scols = [h for h in hdr if h.startswith('stall_') and 'Not' not in h]
    for r in data:
        off = int(r[0], 16) - base
        key = table.get(off, (None, ''))[0]
        e = int(float(r[ix['Instructions Executed']] or 0)) if 'Instructions Executed' in ix else 0
        s = int(float(r[ix['# Samples']] or 0)) if '# Samples' in ix else 0
        ex[key] += e
        sm[key] += s
        for c in scols:
            v = r[ix[c]]
            if v:
                stall[key][c[6:]] += int(float(v))
    te, ts = sum(ex.values()) or 1, sum(sm.values()) or 1
    print(f'kernel {k}: executed {te} samples {ts}')
    for key, s in sm.most_common(top):
        st = ', '.join(f'{a} {b}' for a, b in stall[key].most_common(3))
        print(f'  {str(key):46s} exec {100 * ex[key] / te:5.1f}%  samples {100 * s / ts:5.1f}%  [{st}]')
