#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` dump: stall-reason totals and the hottest SASS lines.
usage: ncu -i rep.ncu-rep --page source --csv --kernel-name regex:X > src.csv; python tools/ncu_hot.py src.csv [top]"""
import csv, sys
from collections import Counter
path = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(open(path)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
col = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = Counter(); lines = []
for n, r in enumerate(rows[hi + 1:]):
    if len(r) < len(hdr): continue
    s = int(r[col["# Samples"]] or 0)
    ex = int(r[col["Instructions Executed"]] or 0)
    st = {h: int(r[col[h]] or 0) for h in stalls}
    for h, v in st.items(): tot[h] += v
    lines.append((s, n, r[col["Source"]].strip(), ex, st, r))
S = sum(l[0] for l in lines)
print("total samples", S, " total warp-instr", sum(l[3] for l in lines))
print("stall totals:", ", ".join(f"{h[6:]}={v*100/S:.1f}%" for h, v in tot.most_common(10)))
print("-- hottest SASS")
for s, n, src, ex, st, r in sorted(lines, reverse=True)[:top]:
    main = ", ".join(f"{h[6:]}:{v}" for h, v in sorted(st.items(), key=lambda kv: -kv[1])[:3] if v)
    extra = ""
    if "L1 Wavefronts Shared" in col and r[col["L1 Wavefronts Shared"]] not in ("", "0"):
        extra = f" shwf={r[col['L1 Wavefronts Shared']]}/ideal {r[col['L1 Wavefronts Shared Ideal']]}"
    print(f"{s*100/S:5.1f}% #{n:4d} ex={ex:>11d} {src[:70]:70s} {main}{extra}")
