#!/usr/bin/env python
"""One-screen summary of an ncu --set full capture (first kernel in the report)."""
import csv, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, v = rows[0], rows[2]
want = ['gpu__time_duration.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'launch__registers_per_thread', 'smsp__warps_active.avg.per_cycle_active',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'dram__bytes_write.sum', 'dram__bytes_read.sum',
        'sm__inst_executed_pipe_alu.sum.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.sum.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.sum.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.sum.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_adu.sum.pct_of_peak_sustained_active', 'sm__icc_request_hit_rate.pct']
for w in want:
    if w in h: print("%-70s %s %s" % (w, v[h.index(w)], rows[1][h.index(w)]))
for i, n in enumerate(h):
    if n.startswith('smsp__average_warps_issue_stalled_') and n.endswith('_per_issue_active.ratio'):
        x = float(v[i])
        if x >= 0.05: print("  stall %-40s %.2f" % (n[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')], x))
