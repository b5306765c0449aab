#!/usr/bin/env python
"""Aggregate ncu source-page stall samples by SASS opcode over the hottest loop.
usage: ncu -i X.ncu-rep --page source --csv > x.csv ; python scripts/stall_by_op.py x.csv [...]"""
import collections
import csv
import sys

COLS = ['stall_dispatch', 'stall_math', 'stall_mio', 'stall_short_sb', 'stall_long_sb', 'stall_wait',
        'stall_not_selected', 'stall_selected', 'stall_barrier']
for path in sys.argv[1:]:
    rows = list(csv.reader(open(path)))
    hdr = rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    body = [r for r in rows[2:] if len(r) > idx['stall_wait']]
    mx = max(int(r[idx['Instructions Executed']] or 0) for r in body)
    agg = collections.defaultdict(lambda: [0] * (len(COLS) + 2))
    tot = 0
    for r in body:
        tot += int(r[idx['# Samples']] or 0)
        if int(r[idx['Instructions Executed']] or 0) >= mx * 0.9:
            toks = r[idx['Source']].split()
            op = toks[1] if toks[0].startswith('@') else toks[0]
            a = agg[op]
            a[0] += 1
            a[1] += int(r[idx['# Samples']] or 0)
            for i, c in enumerate(COLS):
                a[2 + i] += int(r[idx[c]] or 0)
    print(path, 'total samples', tot)
    print('%-14s %4s %7s ' % ('op', 'n', 'samp') + ' '.join(c[6:13].rjust(7) for c in COLS))
    for op, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:14]:
        print('%-14s %4d %7d ' % (op, a[0], a[1]) + ' '.join(str(v).rjust(7) for v in a[2:]))
