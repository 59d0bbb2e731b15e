#!/usr/bin/env python
"""From `ncu -i X.ncu-rep --page raw --csv`: the key metrics of the LAST captured launch of every distinct kernel.
usage: ncu_one_per_kernel.py raw.csv"""
import csv
import sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'l1tex__throughput.avg.pct_of_peak_sustained_active', 'lts__t_sector_hit_rate.pct', 'smsp__inst_executed.sum',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size']
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
last = {}
for r in rows[2:]:
    last[r[hdr.index('Kernel Name')]] = r
for name, r in last.items():
    print('---', name[:90])
    for k in KEYS:
        if k in hdr:
            i = hdr.index(k)
            print(f'  {k} [{units[i]}] = {r[i]}')
