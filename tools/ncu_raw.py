#!/usr/bin/env python
"""Print the handful of `ncu --page raw --csv` columns the profiles/ summaries quote."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]
want = ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'dram__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct', 'launch__registers_per_thread',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct', 'smsp__inst_executed.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_bytes.sum', 'smsp__inst_executed_op_shared_atom.sum',
        'sm__inst_executed_pipe_alu.sum', 'sm__inst_executed_pipe_lsu.sum', 'sm__inst_executed_pipe_fma.sum',
        'smsp__thread_inst_executed_per_inst_executed.ratio']
for w in want:
    if w in hdr:
        i = hdr.index(w)
        print(w, rows[1][i], [r[i][:40] for r in rows[2:]])
for i, h in enumerate(hdr):
    if 'issue_stalled' in h and 'not_issued' not in h and h.endswith('.ratio'):
        vals = [r[i] for r in rows[2:]]
        try:
            if max(float(v) for v in vals) > 0.4:
                print(h.replace('smsp__average_warps_issue_stalled_', '').replace('smsp__average_warp_latency_issue_stalled_', ''), vals)
        except ValueError:
            pass
