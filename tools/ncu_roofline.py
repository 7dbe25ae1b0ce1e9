#!/usr/bin/env python
"""Build profiles/ncu_roofline.json (what bench.py's `roofline` reads) and a text summary from one .ncu-rep that holds
`ncu --set full --clock-control none` captures of the step's kernels at the BENCHMARKED launch shape (a 2048-point
pipeline slice of the scan in scan order).

    python tools/ncu_roofline.py gpurun_out/prof.ncu-rep --points 2048 --tag r02_v2 [--json profiles/ncu_roofline.json]

Per kernel (first captured launch of each name): executed FP64 flops = 2 DFMA + DMUL + DADD thread instructions,
FP64-pipe utilisation, DRAM bytes, duration, occupancy and the stall mix (written to profiles/<tag>_<kernel>_ncu.txt)."""
import argparse
import csv
import json
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEEP = ['gpu__time_duration.sum', 'smsp__cycles_elapsed.avg', 'smsp__sass_thread_inst_executed_op_dfma_pred_on.sum.per_cycle_elapsed', 'smsp__sass_thread_inst_executed_op_dmul_pred_on.sum.per_cycle_elapsed', 'smsp__sass_thread_inst_executed_op_dadd_pred_on.sum.per_cycle_elapsed', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'launch__waves_per_multiprocessor',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fp64.sum', 'smsp__inst_executed.sum',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'smsp__issue_active.avg.per_cycle_active', 'sm__cycles_active.avg', 'sass__inst_executed_local_loads', 'sass__inst_executed_local_stores',
        'sm__inst_executed_pipe_alu.sum', 'sm__inst_executed_pipe_fma.sum', 'sm__inst_executed_pipe_lsu.sum', 'sm__inst_executed_pipe_xu.sum',
        'sm__inst_executed_pipe_cbu.sum', 'sm__inst_executed_pipe_uniform.sum',
        'smsp__sass_thread_inst_executed_op_dfma_pred_on.sum', 'smsp__sass_thread_inst_executed_op_dmul_pred_on.sum', 'smsp__sass_thread_inst_executed_op_dadd_pred_on.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__warps_eligible.avg.per_cycle_active', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed']
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}  # durations -> ms


def fnum(v):
    try:
        return float(v.replace(",", ""))
    except ValueError:
        return None


def short_name(full):
    m = re.match(r"(?:void )?(?:nusi::)?([A-Za-z0-9_]+)(<[^>]*>)?", full)
    name = m.group(1)
    t = m.group(2) or ""
    if name.startswith("rows_") and t:  # the two row kernels differ only in the template argument
        name += "<%s>" % re.sub(r"[^0-9]", "", t.split(",")[0])
    return name


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("--points", type=int, required=True, help="points of the captured launches")
    ap.add_argument("--tag", required=True)
    ap.add_argument("--json", default=os.path.join(ROOT, "profiles", "ncu_roofline.json"))
    ap.add_argument("--command", default="", help="the profiled command line, recorded in the json")
    a = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", a.rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    out = json.load(open(a.json)) if os.path.exists(a.json) else {}
    seen = set()
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        u = dict(zip(hdr, units))
        name = short_name(d.get("Kernel Name", "?"))
        if name in seen:
            continue
        seen.add(name)
        g = lambda k: fnum(d.get(k, ""))
        # the raw page gives the FP64 thread-instruction counts as rates per elapsed cycle (summed over the sub-partitions)
        cyc = g('smsp__cycles_elapsed.avg') or 0.0
        dfma, dmul, dadd = ((g('smsp__sass_thread_inst_executed_op_%s_pred_on.sum' % o) or
                             (g('smsp__sass_thread_inst_executed_op_%s_pred_on.sum.per_cycle_elapsed' % o) or 0.0) * cyc) for o in ("dfma", "dmul", "dadd"))
        dur_ms = g('gpu__time_duration.sum') * UNIT.get(u['gpu__time_duration.sum'], 1.0)
        rd = g('dram__bytes_read.sum') * UNIT.get(u['dram__bytes_read.sum'], 1.0)
        wr = g('dram__bytes_write.sum') * UNIT.get(u['dram__bytes_write.sum'], 1.0)
        out[name] = {
            "points_in_capture": a.points, "duration_ms": dur_ms, "fp64_flop": 2 * dfma + dmul + dadd,
            "fp64_thread_inst": {"dfma": dfma, "dmul": dmul, "dadd": dadd},
            "fp64_pipe_pct": g('sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active'),
            "dram_bytes_read": rd, "dram_bytes_write": wr,
            "registers_per_thread": g('launch__registers_per_thread'), "grid": g('launch__grid_size'), "block": g('launch__block_size'),
            "warps_active_pct": g('sm__warps_active.avg.pct_of_peak_sustained_active'),
            "issue_active_per_cycle": g('smsp__issue_active.avg.per_cycle_active'),
            "warp_inst": g('smsp__inst_executed.sum'),
            "source": "ncu --set full --clock-control none; %s; profiles/%s_*" % (a.command or "bench.py", a.tag),
        }
        lines = ["== %s   (%d points per launch)" % (d.get('Kernel Name', '?'), a.points)]
        for h, uu in zip(hdr, units):
            if h in KEEP or ('issue_stalled' in h and h.endswith('per_issue_active.ratio')):
                v = d[h]
                if 'issue_stalled' in h and (fnum(v) or 0) < 0.05:
                    continue
                lines.append("%s [%s] = %s" % (h, uu, v))
        fn = os.path.join(ROOT, "profiles", "%s_%s_ncu.txt" % (a.tag, re.sub(r"[<>]", "", name)))
        open(fn, "w").write("\n".join(lines) + "\n")
        print("%-28s %8.3f ms  fp64 pipe %5.1f %%  %7.2f MFLOP/point  regs %3d  dram rd %.1f MB wr %.1f MB" % (
            name, dur_ms, out[name]["fp64_pipe_pct"] or 0, out[name]["fp64_flop"] / a.points / 1e6, out[name]["registers_per_thread"] or 0, rd / 1e6, wr / 1e6))
    json.dump(out, open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
