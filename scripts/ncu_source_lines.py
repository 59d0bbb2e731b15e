#!/usr/bin/env python
"""Per CUDA source line: warp instructions and stall samples from `ncu --page source --csv --print-source cuda,sass`."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file = ''
hdr = None
lines = []
for r in rows:
    if not r:
        continue
    if r[0] == 'File Path':
        cur_file = r[1].split('/')[-1]
    elif r[0] == 'Line No':
        hdr = r
        ix_inst = hdr.index('Instructions Executed'); ix_samp = hdr.index('# Samples')
    elif hdr and r[0] not in ('', 'Function Name') and r[2] == '-':
        try:
            lines.append((int(r[ix_inst] or 0), int(r[ix_samp] or 0), cur_file, r[0], r[1].strip()[:110]))
        except ValueError:
            pass
ti = sum(l[0] for l in lines); ts = sum(l[1] for l in lines)
print(f'total warp inst {ti}, samples {ts}')
for inst, samp, f, ln, src in sorted(lines, key=lambda x: -x[0])[:top]:
    print(f'{100*inst/ti:5.1f}% inst {100*samp/max(ts,1):5.1f}% smp  {f}:{ln:>4s}  {src}')
