#!/usr/bin/env python
"""profiles/ncu_roofline.json from the --set full captures of the hot kernels (what bench.py quotes as
roofline.fp64_pipe_pct / roofline.traffic):  python tools/ncu_roofline_json.py TAG name=rep [name=rep ...]"""
import csv, json, subprocess, sys
tag, out = sys.argv[1], {}
MULT = {'Mbyte': 1e6, 'Kbyte': 1e3, 'Gbyte': 1e9, 'byte': 1, 'ms': 1e-3, 'us': 1e-6, 'ns': 1e-9, 's': 1}
for arg in sys.argv[2:]:
    name, rep = arg.split("=")
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    d, un = dict(zip(rows[0], rows[2])), dict(zip(rows[0], rows[1]))
    val = lambda k: float(d[k].replace(',', '')) * MULT.get(un[k], 1)
    out[name] = {"points_in_capture": int(float(d['launch__grid_size'].replace(',', ''))) if name.startswith("integrate") else 296,
                 "seconds": val('gpu__time_duration.sum'), "dram_bytes_read": val('dram__bytes_read.sum'),
                 "dram_bytes_write": val('dram__bytes_write.sum'),
                 "fp64_pipe_pct": float(d['sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active']),
                 "issue_active_per_cycle": float(d['smsp__issue_active.avg.per_cycle_active']),
                 "registers_per_thread": int(float(d['launch__registers_per_thread'])),
                 "warp_inst": float(d['smsp__inst_executed.sum'].replace(',', '')),
                 # FP64 flops the launch executed: (2 DFMA + DMUL + DADD) thread instructions per elapsed cycle x elapsed cycles
                 "fp64_flop": (2 * float(d['smsp__sass_thread_inst_executed_op_dfma_pred_on.sum.per_cycle_elapsed'].replace(',', ''))
                               + float(d['smsp__sass_thread_inst_executed_op_dmul_pred_on.sum.per_cycle_elapsed'].replace(',', ''))
                               + float(d['smsp__sass_thread_inst_executed_op_dadd_pred_on.sum.per_cycle_elapsed'].replace(',', '')))
                 * float(d['smsp__cycles_elapsed.avg'].replace(',', '')),
                 "source": "ncu --set full --clock-control none, bench.py --points 296 --steps 2 --warmup 1 (profiles/%s_*)" % tag}
json.dump(out, open("profiles/ncu_roofline.json", "w"), indent=1)
print(json.dumps(out, indent=1))
