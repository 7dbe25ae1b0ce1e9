#!/usr/bin/env python
"""Summarise one kernel of an .ncu-rep (raw page) into a short text file for profiles/."""
import csv, subprocess, sys
rep, out = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else None)
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
keep = ['Kernel Name', 'gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fp64.sum', 'smsp__inst_executed.sum',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'smsp__issue_active.avg.per_cycle_active', 'sm__cycles_active.avg', 'sass__inst_executed_local_loads', 'sass__inst_executed_local_stores',
        'sm__inst_executed_pipe_alu.sum', 'sm__inst_executed_pipe_fma.sum', 'sm__inst_executed_pipe_lsu.sum', 'sm__inst_executed_pipe_xu.sum',
        'sm__inst_executed_pipe_fmaheavy.sum', 'sm__inst_executed_pipe_cbu.sum', 'sm__inst_executed_pipe_uniform.sum',
        'smsp__sass_thread_inst_executed_op_dfma_pred_on.sum', 'smsp__sass_thread_inst_executed_op_dmul_pred_on.sum', 'smsp__sass_thread_inst_executed_op_dadd_pred_on.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__warps_eligible.avg.per_cycle_active']
lines = []
for r in rows[2:]:
    d = dict(zip(hdr, r))
    lines.append("== %s" % d.get('Kernel Name', '?'))
    for h, u in zip(hdr, units):
        if h in keep[1:] or ('issue_stalled' in h and h.endswith('per_issue_active.ratio')):
            v = d[h]
            if 'issue_stalled' in h:
                try:
                    if float(v) < 0.05: continue
                except ValueError: pass
            lines.append("%s [%s] = %s" % (h, u, v))
txt = "\n".join(lines)
print(txt)
if out: open(out, "a").write(txt + "\n")
