#!/usr/bin/env python
"""Group the SASS of an `ncu --page source --csv` dump into runs of equal execution count (~ basic blocks of one loop
level) and print each run's share of instructions and stall samples.  usage: python tools/ncu_blocks.py src.csv [min_pct]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1]))); minp = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
hdr = rows[hi]; col = {h: i for i, h in enumerate(hdr)}
L = []
for n, r in enumerate(rows[hi + 1:]):
    if len(r) < len(hdr): continue
    L.append((n, int(r[col['Instructions Executed']] or 0), int(r[col['# Samples']] or 0), r[col['Source']].strip()))
tot = sum(x[1] for x in L) or 1; ts = sum(x[2] for x in L) or 1
runs = []
for n, ex, s, src in L:
    if runs and runs[-1]['ex'] and abs(ex - runs[-1]['ex']) <= 0.02 * runs[-1]['ex']:
        r = runs[-1]; r['end'] = n; r['inst'] += ex; r['smp'] += s; r['ops'].append(src.split()[0] if src else '')
    else:
        runs.append({'start': n, 'end': n, 'ex': ex, 'inst': ex, 'smp': s, 'ops': [src.split()[0] if src else '']})
print('total warp-instr', tot, 'samples', ts)
for r in runs:
    if r['inst'] * 100 / tot >= minp or r['smp'] * 100 / ts >= minp:
        ops = ' '.join(r['ops'][:14])
        print(f"#{r['start']:4d}-{r['end']:4d} n={r['end']-r['start']+1:3d} ex={r['ex']:>10d} inst {r['inst']*100/tot:5.1f}% smp {r['smp']*100/ts:5.1f}%  {ops}")
