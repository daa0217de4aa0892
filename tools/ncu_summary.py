#!/usr/bin/env python
"""Summarise `ncu --page raw --csv` / `--page source --csv` exports (used to write profiles/*.txt)."""
import csv
import sys

KEEP = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__inst_executed.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__sass_inst_executed_op_local_ld.sum', 'smsp__sass_inst_executed_op_local_st.sum',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
        'sm__cycles_active.avg', 'smsp__cycles_active.avg', 'sm__cycles_elapsed.max', 'smsp__warps_eligible.avg.per_cycle_active',
        'smsp__warps_active.avg.per_cycle_active', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_bytes.sum', 'l1tex__t_bytes.sum']


def raw(path):
    rows = list(csv.reader(open(path)))
    for h, u, v in zip(rows[0], rows[1], rows[2]):
        if h in KEEP or h == 'Kernel Name':
            print(f'{h:90s} {v:>24s} {u}')


def source(path, top=30):
    rows = list(csv.reader(open(path)))
    hdr, data = rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    tot_s = sum(int(r[ix['# Samples']]) for r in data)
    tot_i = sum(int(r[ix['Instructions Executed']]) for r in data)
    print(f'-- source page: {len(data)} SASS instructions, {tot_i/1e9:.2f} G warp-instructions executed, {tot_s} stall samples; top {top} by samples')
    for r in sorted(data, key=lambda r: -int(r[ix['# Samples']]))[:top]:
        print(f"{int(r[ix['# Samples']])/tot_s*100:5.2f}% inst={int(r[ix['Instructions Executed']])/1e6:8.1f}M thr={r[ix['Avg. Threads Executed']]:>5s} {r[ix['Source']].strip()[:90]}")


if __name__ == '__main__':
    raw(sys.argv[1])
    if len(sys.argv) > 2:
        source(sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 30)
