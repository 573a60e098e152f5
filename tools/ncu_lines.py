#!/usr/bin/env python
"""Per-CUDA-source-line instruction and sample shares from `ncu -i rep --page source --csv --print-source cuda`.
usage: python tools/ncu_lines.py src_cuda.csv [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1]))); top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hi = next(i for i, r in enumerate(rows) if r and 'Source' in r and 'Instructions Executed' in r)
hdr = rows[hi]; col = {h: i for i, h in enumerate(hdr)}
out = []
for r in rows[hi + 1:]:
    if len(r) < len(hdr): continue
    try:
        ex = int(r[col['Instructions Executed']] or 0); s = int(r[col['# Samples']] or 0)
    except ValueError:
        continue
    out.append((ex, s, r[0], r[col['Source']].strip()[:120]))
tot = sum(o[0] for o in out) or 1; ts = sum(o[1] for o in out) or 1
print('total warp-instr', tot, 'samples', ts)
for ex, s, n, src in sorted(out, reverse=True)[:top]:
    print(f"{ex*100/tot:5.1f}% inst {s*100/ts:5.1f}% smp  L{n}: {src}")
