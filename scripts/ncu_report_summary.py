#!/usr/bin/env python
"""Summarises `ncu -i X.ncu-rep --page raw --csv` (arg 1) and optionally `--page source --csv` (arg 2)."""
import collections
import csv
import re
import sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_warps', 'launch__occupancy_limit_blocks',
        'launch__waves_per_multiprocessor', 'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__inst_executed.sum',
        'launch__shared_mem_per_block_dynamic', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'launch__block_size', 'launch__grid_size', 'sm__cycles_elapsed.max', 'lts__t_sector_hit_rate.pct',
        'l1tex__t_sector_hit_rate.pct', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'sm__inst_executed_pipe_fma.sum', 'sm__inst_executed_pipe_alu.sum', 'sm__inst_executed_pipe_lsu.sum',
        'l1tex__data_pipe_lsu_wavefronts.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'lts__t_bytes.sum', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
        'l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_active', 'l1tex__throughput.avg.pct_of_peak_sustained_active']
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    print('---', r[hdr.index('Kernel Name')][:80])
    for k in KEYS:
        if k in hdr:
            i = hdr.index(k)
            print(f'  {k} [{units[i]}] = {r[i]}')
if len(sys.argv) > 2:
    rows = list(csv.reader(open(sys.argv[2])))
    hi = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
    start = hi[0]
    end = hi[1] - 1 if len(hi) > 1 else len(rows)
    hdr = rows[start]
    ix = {k: i for i, k in enumerate(hdr)}
    body = [r for r in rows[start + 1:end] if len(r) > 10]
    tot_inst = sum(int(r[ix['Instructions Executed']] or 0) for r in body)
    tot_samp = sum(int(r[ix['# Samples']] or 0) for r in body)
    print(f'source page: {len(body)} SASS lines, {tot_inst} warp instructions, {tot_samp} samples')
    stalls = [k for k in hdr if k.startswith('stall_') and 'Not Issued' not in k]
    agg = {k: sum(int(r[ix[k]] or 0) for r in body) for k in stalls}
    print('  stalls: ' + ', '.join(f'{k[6:]} {100 * v / max(tot_samp, 1):.1f}%' for k, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
    op = collections.Counter()
    for r in body:
        m = re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)', r[ix['Source']])
        if m:
            op[m.group(2).split('.')[0]] += int(r[ix['Instructions Executed']] or 0)
    print('  opcodes: ' + ', '.join(f'{k} {100 * v / tot_inst:.1f}%' for k, v in op.most_common(16)))
