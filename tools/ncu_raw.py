#!/usr/bin/env python
"""Print the key metrics of every kernel in an `ncu --page raw --csv` dump.  usage: tools/ncu_raw.py raw.csv"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
h = rows[0]
ix = {k: i for i, k in enumerate(h)}
want = ['gpu__time_duration.sum', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'smsp__inst_executed.sum', 'launch__registers_per_thread',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'sm__inst_executed_pipe_tensor.sum', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active']
for r in rows[2:]:
    print('----', r[ix['Kernel Name']][:70])
    for w in want:
        if w in ix:
            print(f'  {w:72s} {r[ix[w]]} {rows[1][ix[w]]}')
    for k in h:
        if k.startswith('smsp__average_warps_issue_stalled') and k.endswith('per_issue_active.ratio'):
            v = float(r[ix[k]])
            if v > 0.1:
                print(f'  stall {k[34:-23]:40s} {v:.2f}')
