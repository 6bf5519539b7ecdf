#!/usr/bin/env python3
"""Summarise `ncu -i X.ncu-rep --page raw --csv` output: the handful of per-kernel metrics quoted in DESIGN.md / bench.py.
usage: ncu_raw_summary.py raw.csv"""
import csv
import sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__grid_size', 'launch__block_size',
        'launch__registers_per_thread', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed',
        'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__average_warp_latency_per_inst_issued.ratio']
rows = list(csv.reader(open(sys.argv[1])))
h, units = rows[0], rows[1]
for r in rows[2:]:
    print('==', r[h.index('Kernel Name')])
    for k in KEYS:
        if k in h:
            print(f'  {k:82s} {r[h.index(k)]:>16s} {units[h.index(k)]}')
    st = [(k.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''), float(r[i] or 0))
          for i, k in enumerate(h) if k.startswith('smsp__average_warps_issue_stalled_') and k.endswith('_per_issue_active.ratio')]
    print('  warps stalled per issue-active cycle:', ', '.join(f'{k} {v:.2f}' for k, v in sorted(st, key=lambda kv: -kv[1]) if v >= 0.05))
