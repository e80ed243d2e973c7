#!/usr/bin/env python
"""Merge `ncu -i X.ncu-rep --page raw --csv` dumps into one table of the columns the profile summaries quote,
durations normalised to us and byte counts to Mbyte (captures differ in the units ncu picks per column).
usage: tools/ncu_merge.py out.csv raw_a.csv raw_b.csv ..."""
import csv
import sys

COLS = ['gpu__time_duration.sum', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'smsp__inst_executed.sum', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__block_size', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'lts__t_sector_hit_rate.pct',
        'smsp__average_warp_latency_issue_stalled_no_instruction.ratio',
        'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio']
TIME = {'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 's': 1e6}
BYTES = {'byte': 1e-6, 'Kbyte': 1e-3, 'Mbyte': 1.0, 'Gbyte': 1e3}


def main():
    out, srcs = sys.argv[1], sys.argv[2:]
    table, units_out, cols = [], {}, None
    for src in srcs:
        rows = list(csv.reader(open(src)))
        h, units = rows[0], rows[1]
        ix = {k: i for i, k in enumerate(h)}
        have = [c for c in COLS if c in ix]
        cols = have if cols is None else [c for c in cols if c in have]
        for r in rows[2:]:
            rec = {'Kernel Name': r[ix['Kernel Name']]}
            for c in have:
                v, u = r[ix[c]].replace(',', ''), units[ix[c]]
                if u in TIME and c.startswith('gpu__time'):
                    v, u = '%.3f' % (float(v) * TIME[u]), 'us'
                elif u in BYTES:
                    v, u = '%.3f' % (float(v) * BYTES[u]), 'Mbyte'
                rec[c] = v
                units_out[c] = u
            table.append(rec)
    with open(out, 'w', newline='') as f:
        w = csv.writer(f)
        w.writerow(['Kernel Name'] + cols)
        w.writerow([''] + [units_out[c] for c in cols])
        for rec in table:
            w.writerow([rec['Kernel Name']] + [rec[c] for c in cols])
    print(f'{out}: {len(table)} launches, {len(cols)} columns')


if __name__ == '__main__':
    main()
