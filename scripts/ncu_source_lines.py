#!/usr/bin/env python
"""Per CUDA source line: warp instructions and stall samples from
`ncu -i X.ncu-rep --page source --csv --print-source cuda,sass`.   usage: ncu_source_lines.py file.csv [top]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur, hdr, agg = None, None, {}
for r in rows:
    if not r:
        continue
    if r[0] == 'File Path':
        cur = r[1].split('/')[-1]
    elif r[0] == 'Line No':
        hdr = r
        ii, isamp, ith = hdr.index('Instructions Executed'), hdr.index('# Samples'), hdr.index('Thread Instructions Executed')
    elif hdr and len(r) > ii and r[0] != '' and r[2] == '-':
        try:
            agg[(cur, int(r[0]))] = (int(r[ii] or 0), int(r[isamp] or 0), int(r[ith] or 0), r[1].strip()[:100])
        except ValueError:
            pass
ti = sum(v[0] for v in agg.values())
ts = sum(v[1] for v in agg.values())
print(f'total warp inst {ti}, samples {ts}')
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f'{100 * v[0] / ti:5.1f}% inst {100 * v[1] / max(ts, 1):5.1f}% smp  thr/inst {v[2] / max(v[0], 1):4.1f}  {k[0]}:{k[1]:4d}  {v[3]}')
