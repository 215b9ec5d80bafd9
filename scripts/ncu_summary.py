#!/usr/bin/env python
"""Key counters of an ncu report (raw page): usage ncu_summary.py report.ncu-rep"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h = rows[0]
for v in rows[2:]:
    print("kernel:", v[h.index("Kernel Name")][:90])
    want = ['gpu__time_duration.sum', 'smsp__inst_executed.sum', 'sm__inst_issued.avg.pct_of_peak_sustained_active',
            'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'dram__bytes_read.sum',
            'dram__bytes_write.sum', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'launch__grid_size',
            'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_warps',
            'l1tex__data_pipe_lsu_wavefronts.sum', 'smsp__inst_executed_pipe_lsu.sum',
            'l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_active', 'l1tex__data_bank_conflicts_pipe_lsu.sum',
            'sm__inst_executed_pipe_alu.sum', 'lts__t_bytes.sum', 'l1tex__throughput.avg.pct_of_peak_sustained_active',
            'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'dram__throughput.avg.pct_of_peak_sustained_elapsed']
    for w in want:
        if w in h:
            print(f"  {w} = {v[h.index(w)]} {rows[1][h.index(w)]}")
    for i, n in enumerate(h):
        if 'issue_stalled' in n and 'per_issue_active' in n:
            try:
                if float(v[i]) > 0.25:
                    print("  stall", n.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''), v[i])
            except ValueError:
                pass
