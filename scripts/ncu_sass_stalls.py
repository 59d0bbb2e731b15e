#!/usr/bin/env python
"""Top stalled SASS instructions of `ncu -i X.ncu-rep --page source --csv` (arg 1), with their dominant stall reasons."""
import csv
import sys
rows = list(csv.reader(open(sys.argv[1])))
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
hi = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
hdr = rows[hi[0]]
ix = {k: i for i, k in enumerate(hdr)}
body = [r for r in rows[hi[0] + 1:] if len(r) > 10]
tot = sum(int(r[ix['# Samples']] or 0) for r in body)
toti = sum(int(r[ix['Instructions Executed']] or 0) for r in body)
print(f'{len(body)} SASS lines, {toti} warp instructions, {tot} samples')
for r in sorted(body, key=lambda r: -int(r[ix['# Samples']] or 0))[:top_n]:
    st = {k[6:]: int(r[ix[k]] or 0) for k in hdr if k.startswith('stall_') and 'Not Issued' not in k}
    best = sorted(st.items(), key=lambda x: -x[1])[:2]
    print(f"{100 * int(r[ix['# Samples']]) / tot:5.2f}%  inst={r[ix['Instructions Executed']]:>9s}  {r[ix['Source']][:64]:64s} {best}")
