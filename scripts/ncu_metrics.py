#!/usr/bin/env python
"""Key metrics and top stall reasons per kernel of an `ncu --page raw --csv` export.  usage: python scripts/ncu_metrics.py gpurun_out/kp_raw.csv"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
want = ['gpu__time_duration.sum', 'sm__cycles_elapsed.max', 'smsp__inst_executed.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.avg.per_cycle_active', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__registers_per_thread',
        'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
        'sm__maximum_warps_per_active_cycle_pct', 'launch__grid_size', 'launch__block_size']
stalls = [h for h in hdr if h.startswith('smsp__average_warps_issue_stalled_') and h.endswith('_per_issue_active.ratio')]
if not stalls:
    stalls = [h for h in hdr if 'issue_stalled' in h]
for r in rows[2:]:
    print('==', r[idx['Kernel Name']][:100])
    for w in want:
        if w in idx:
            print(f"   {w:70s} {r[idx[w]]:>16s} {units[idx[w]]}")
    vals = []
    for h in stalls:
        try:
            vals.append((float(r[idx[h]].replace(',', '') or 0), h))
        except ValueError:
            pass
    for v, h in sorted(vals, reverse=True)[:7]:
        print(f"   stall {v:9.3f}  {h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')}")
