#!/usr/bin/env python
"""bench.py -- the reference's headline metric on B200: (mphi, g) evolutions per second at N_E = 200.

Workload (BASELINE.json configs[4], SURVEY 8(d) "S-scan"): the 256 x 256 parameter scan
    mphi_i = 10^(1 + 3 i/255) eV,  g_j = 10^(-5 + 5 j/255),  mntot = 0.1 eV,
    Majorana, non_resonant, normal ordering, N_bins_E = 200, lE 4 -> 9, zmax = 5, flav = 2, phiphi off
sharded statically over the 8 GPUs of a box: scan point p belongs to shard p mod 8.  One "step" is
one evolve() of every point of a shard (8192 points) per GPU: with --gpus 8 a step is the whole
scan, with fewer GPUs it is the first N shards (weak scaling, fixed per-GPU work).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...   # the reference's own CPU evolve()
    python bench.py --workload fine [--points P]              # BASELINE configs[3]: N_E = 2000 (not the headline line)

One JSON line on stdout (rank 0).  `value` times the device-resident path of the product's scan API
(nusiprop_b200.scan.ScanRunner: inputs already in HBM, outputs left in HBM, gather to rank 0 with N > 1); `e2e`
times nusi_evolve_batch through the C ABI with HOST buffers (H2D of the parameters and D2H of both flux arrays
inside the timed region).  At N = 1 the line also carries `parity` -- the whole scan evolved in the same run and
compared with the oracle fixture of the seeded 1024-point subsample (tests/scan_parity.py) -- `roofline` (executed
FP64 flops of each kernel / the DFMA peak measured in the same run) and `cpu_baseline`.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np

N_E, LEMIN, LEMAX, ZMAX, FLAV = 200, 4.0, 9.0, 5.0, 2
MNTOT = 0.1
N_SHARDS = 8
# algorithmic FP64 work (add/mul/div = 1, FMA = 2, log/exp = 32, atan = 48, Li2 = 96,
# complex Li2 = 320) after common-subexpression elimination
KFLOP_ALPHA_ENTRY = 13.9e3
KFLOP_ROW = 22.4e3
FLOP_LUMINT = 308.0


def scan_points():
    i, j = np.meshgrid(np.arange(256), np.arange(256), indexing="ij")
    mphi = 10.0 ** (1.0 + 3.0 * i / 255.0)
    g = 10.0 ** (-5.0 + 5.0 * j / 255.0)
    return mphi.ravel(), g.ravel()


def shard(rank):
    mphi, g = scan_points()
    idx = np.arange(rank, mphi.size, N_SHARDS)
    return np.ascontiguousarray(mphi[idx]), np.ascontiguousarray(g[idx])


def flops_per_evolution(NT, N, Nz):
    table = NT * (NT - 1) / 2 * KFLOP_ALPHA_ENTRY
    rows = NT * KFLOP_ROW
    integ = (Nz - 1) * (N * (N - 1) + 90.0 * N)
    return table, rows, integ


class ClockSampler(object):
    """nvidia-smi clock / throttle-reason samples DURING the timed region (profiling recipe)."""

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------
# CPU reference arm: the reference's own evolve() (oracle/_ref = unmodified headers + refshim),
# one process per parameter point, all host cores
# ---------------------------------------------------------------------------------------------
def ref_binary():
    for name in ("nusi_ref_o3", "nusi_ref"):
        p = os.path.join(ROOT, "oracle", "_ref", name)
        if os.path.exists(p):
            return p
    return None


def run_reference_sample(points, cores, n_e=N_E):
    """Run evolve() for `points` [(mphi, g)] with `cores` concurrent single-point processes.
    Returns wall seconds."""
    exe = ref_binary()
    if exe is None:
        raise RuntimeError("oracle/_ref/nusi_ref_o3 missing (build it with `make -C oracle` where /root/reference exists)")
    it = iter(points)
    lock = threading.Lock()

    def worker():
        while True:
            with lock:
                try:
                    mphi, g = next(it)
                except StopIteration:
                    return
            subprocess.run([exe, "evolve", repr(float(mphi)), repr(float(g)), repr(MNTOT), "2", "1", "1", "1", "1", str(n_e),
                            repr(LEMIN), repr(LEMAX), repr(ZMAX), str(FLAV), "0", "1", "1"],
                           stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, check=True)

    t0 = time.perf_counter()
    th = [threading.Thread(target=worker) for _ in range(cores)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    return time.perf_counter() - t0


def cpu_sample_points(count):
    """Seeded subsample of the scan (default_rng(20261104).choice(65536, 1024))."""
    mphi, g = scan_points()
    idx = np.random.default_rng(20261104).choice(mphi.size, 1024, replace=False)[:count]
    return [(mphi[i], g[i]) for i in idx]


FINE_N_E = 2000
FINE_POINTS = [(100.0, 1e-2), (100.0, 1e-1), (30.0, 1e-2), (300.0, 1e-1)]  # first one = BASELINE configs[3] / SURVEY S-fine


def bench_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = os.cpu_count() or 1
    fine = args.workload == "fine"
    if fine:  # one N_E = 2000 evolve() takes ~34 s on one core: one point per core per step
        per_step = min(cores, len(FINE_POINTS))
        pts = lambda cnt: (FINE_POINTS * ((cnt + len(FINE_POINTS) - 1) // len(FINE_POINTS)))[:cnt]
        n_e = FINE_N_E
        for _ in range(min(args.warmup, 1)):
            run_reference_sample(pts(per_step), cores, n_e)
    else:
        per_step = max(cores * 2, 16)  # ~2 evolve() per core per step, ~0.35 s each
        pts = cpu_sample_points
        n_e = N_E
        for _ in range(args.warmup):
            run_reference_sample(pts(min(cores, 16)), cores, n_e)
    t = 0.0
    steps = min(args.steps, 2) if fine else args.steps
    for s in range(steps):
        t += run_reference_sample(pts(per_step), cores, n_e)
    value = per_step * steps / t
    sample = "%d evolve() per step (%s), one process per point, %d concurrent" % (
        per_step, "N_E=2000 points" if fine else "seeded subsample of the 256x256 scan", cores)
    line = {
        "metric": "(mphi,g) evolutions/sec at N_E=%d" % n_e, "value": value, "unit": "evolutions/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "impl": "reference",
        "config": fine_config(per_step) if fine else workload_config(args.gpus),
        "cpu_baseline": {"value": value, "unit": "evolutions/s", "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": "evolutions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def fine_config(points):
    return {"workload": "S-fine (BASELINE configs[3]): N_E=2000, lE 4->9 (N_z=313, NT=2311), zmax=5, Majorana, non_resonant, NO, flav=2, "
                        "phiphi off, mntot=0.1; %d point(s) per step" % points,
            "points_per_gpu_per_step": points, "N_bins_E": FINE_N_E,
            "l2": "per-point alpha table 21.4 MB; the batch's tables exceed L2 from 6 points on"}


def workload_config(gpus):
    return {"workload": "S-scan: 256x256 (mphi,g) grid, shard p mod 8 per GPU (8192 points per GPU per step), N_E=200, lE 4->9, "
                        "zmax=5, Majorana, non_resonant, NO, flav=2, phiphi off, mntot=0.1",
            "points_per_gpu_per_step": 8192, "N_bins_E": N_E, "parallelism": "points sharded over %d GPU(s), gather to rank 0" % gpus,
            "l2": "per-step working set (packed alpha tables, 1.7 GB per GPU) exceeds the 126 MB L2; no flush needed"}


def roofline(peak, points, steps, ms_kernel, pipeline_slice=2048):
    """Each kernel of the step against the FP64 pipe -- the binding resource (arithmetic intensity ~4e4
    flop/B against a ridge of ~6; no contraction anywhere, so tensor cores have nothing to do).
      achieved  = FP64 flops the kernel EXECUTES per point (2 x DFMA + DMUL + DADD thread instructions of the committed
                  ncu capture of the same launch shape, profiles/ncu_roofline.json) x points / CUDA-event time measured
                  live in this run;   peak = DFMA rate measured in this run (MEASURED_PEAKS.json has no FP64 figure);
      frac      = achieved / peak.  fp64_pipe_pct = ncu sm__pipe_fp64_cycles_active of the capture (pipe utilisation:
                  higher than frac because only part of the FP64 instructions are FMAs -- contraction is off for parity).
      reference_equivalent_x = SURVEY 8(d)'s per-entry census of the REFERENCE's formulation (13.9 kflop per table entry)
                  / the same time / peak: how much reference work per second the step replaces, NOT a utilisation
                  (the kernels share vertex work between entries and g-independent work between the points of a scan row).
      traffic   = ncu dram bytes read + written per launch of one 2048-point pipeline slice."""
    out = {"bound": "fp64", "kernel": "alpha_group_kernel", "achieved": None, "peak": peak, "unit": "TFLOP/s", "frac": None,
           "traffic": None, "peak_source": "DFMA rate measured in this run by nusi_measure_fp64_peak (8 independent chains per thread)"}
    try:
        prof = json.load(open(os.path.join(ROOT, "profiles", "ncu_roofline.json")))
    except Exception as ex:
        out["note"] = "profiles/ncu_roofline.json unavailable: %s" % ex
        return out, 0.0, 0.0, 0.0
    # the default integrator at N_E <= 224 is integrate_split_kernel (integrate_diag_kernel before r02_v3 and for opts.integrator = 3)
    integ = "integrate_split_kernel" if "integrate_split_kernel" in prof else "integrate_diag_kernel"
    edges = "edges_group_kernel" if "edges_group_kernel" in prof else "edges_kernel"  # the scan builder's edge kernel since r02_v11
    fam = {"alpha_kernel": ["alpha_group_kernel", edges], "integrate_kernel": [integ],
           "rows_kernel": ([n for n in ("rows_shared_kernel<0>", "rows_point_kernel<0>", "rows_shared_kernel<1>", "rows_point_kernel<1>") if n in prof]
                           or ["rows_group_kernel<0>", "rows_group_kernel<1>"])}  # two kernels per function since r02_v12
    kernels, tot_flop, tot_busy, tot_ms = {}, 0.0, 0.0, 0.0
    for key, names in fam.items():
        ms = ms_kernel.get(key, 0.0)
        for nm in names:
            k = prof.get(nm)
            if not k:
                continue
            flop_pp = k["fp64_flop"] / k["points_in_capture"]
            share = k["duration_ms"] / sum(prof[n]["duration_ms"] for n in names if n in prof)  # this kernel's share of the family's event time
            t = ms * share
            e = {"ms_per_step": t / steps, "executed_mflop_per_point": flop_pp / 1e6, "fp64_pipe_pct": k["fp64_pipe_pct"],
                 "achieved": flop_pp * points * steps / (t * 1e-3) / 1e12 if t > 0 else None,
                 "dram_bytes_per_slice": (k["dram_bytes_read"] + k["dram_bytes_write"]) * pipeline_slice / k["points_in_capture"],
                 "ncu_source": k["source"]}
            e["frac"] = e["achieved"] / peak if e["achieved"] else None
            kernels[nm] = e
            tot_flop += flop_pp * points * steps
            tot_busy += t * k["fp64_pipe_pct"] / 100.0
            tot_ms += t
    out["kernels"] = kernels
    d = kernels.get("alpha_group_kernel")
    if d:
        out.update(achieved=d["achieved"], frac=d["frac"], traffic=d["dram_bytes_per_slice"], fp64_pipe_pct=d["fp64_pipe_pct"],
                   ncu_source=d["ncu_source"])
    return out, tot_flop, tot_busy, tot_ms


def validate_outputs(fm, ff, st, what):
    """The timed workload must be a real result: every status word clean up to the reference's own warnings, every flux finite."""
    bad = int(np.sum((st & ~7) != 0))
    if bad or not np.isfinite(fm).all() or not np.isfinite(ff).all():
        raise RuntimeError("%s: %d points with a fatal status, finite mass %s, finite flavour %s" % (what, bad, np.isfinite(fm).all(), np.isfinite(ff).all()))
    return {"points": int(st.size), "status_clean": int(np.sum(st == 0)), "status_warn_only": int(np.sum((st != 0) & ((st & ~7) == 0))), "all_finite": True}


# ---------------------------------------------------------------------------------------------
# CUDA arm, headline workload
# ---------------------------------------------------------------------------------------------
def bench_cuda(args):
    import torch
    import torch.distributed as dist
    import nusiprop_b200 as nb
    from nusiprop_b200.scan import ScanRunner

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)

    ctx = nb.Context(majorana=True, non_resonant=True, normal_ordering=True, N_bins_E=N_E, lEmin=LEMIN, lEmax=LEMAX, zmax=ZMAX,
                     flav=FLAV, phiphi=False, device=local)
    # weak scaling: N GPUs evolve the first N shards of the scan, one 8192-point shard each.  The product's scan API
    # shards a point list p -> rank p mod world, so hand it the interleaved union of those shards.
    mphi_all, g_all = scan_points()
    sel = np.arange(mphi_all.size).reshape(-1, N_SHARDS)[:, :world].ravel()  # points of shards 0..world-1, interleaved
    if args.points:
        sel = sel[: args.points * world]
    run = ScanRunner(ctx, mphi_all[sel], g_all[sel], MNTOT)
    n = run.n_local
    mphi, g = mphi_all[sel][run.idx], g_all[sel][run.idx]
    assert np.array_equal(mphi, shard(rank % N_SHARDS)[0][:n])
    mntot = np.full(n, MNTOT)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 0)):
        run.step()
    barrier()
    ctx.last_timing()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        run.step()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    tm = ctx.last_timing()  # per-kernel CUDA-event sums over the timed steps
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    clocks = sampler.stop() if rank == 0 else None
    res = run.result()
    d_fm, d_ff, d_st = run.flux_mass[:n].cpu().numpy(), run.flux_fla[:n].cpu().numpy(), run.status[:n].cpu().numpy()
    check = validate_outputs(d_fm, d_ff, d_st, "device-resident path, rank %d" % rank)
    if world > 1 and rank == 0:  # the gathered scan is what a consumer reads: same numbers as each rank's own
        assert np.array_equal(res[1][rank::world], d_ff)
        validate_outputs(res[1], res[1], res[2], "gathered scan")

    # end-to-end through the C ABI with host buffers (pinned), every rank concurrently
    h_fm = torch.empty((n, 3, N_E), dtype=torch.float64, pin_memory=True).numpy()
    h_ff = torch.empty((n, 3, N_E), dtype=torch.float64, pin_memory=True).numpy()
    from nusiprop_b200.api import _dp, _ip
    h_st = torch.zeros(n, dtype=torch.int32, pin_memory=True).numpy()
    lib = ctx._lib

    def e2e_step():
        rc = lib.nusi_evolve_batch(ctx._h, n, _dp(mphi), _dp(g), _dp(mntot), _dp(h_fm), _dp(h_ff), _ip(h_st))
        if rc != 0:
            raise RuntimeError("nusi_evolve_batch rc=%d" % rc)

    e2e_steps = max(1, min(args.steps, 3))
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    # the device path and the host path must agree bit for bit
    same = bool(np.array_equal(d_ff, h_ff, equal_nan=True) and np.array_equal(d_fm, h_fm, equal_nan=True) and np.array_equal(d_st, h_st))
    if not same:
        raise RuntimeError("host-buffer path and device-resident path disagree")

    if rank == 0:
        N, Nz, NT = ctx.N_bins_E, ctx.N_steps_z, ctx.NT
        f_tab, f_rows, f_int = flops_per_evolution(NT, N, Nz)
        total_ev = n * world * args.steps
        value = total_ev / (ms * 1e-3)
        peak = nb.measure_fp64_peak(local)
        ms_kernel = {"rows_kernel": tm["ms_rows"], "alpha_kernel": tm["ms_alpha"], "integrate_kernel": tm["ms_integrate"]}
        rf, tot_flop, tot_busy, tot_ms = roofline(peak, n, args.steps, ms_kernel)
        if tot_ms > 0:
            rf["step"] = {"executed_tflops": tot_flop / (ms * 1e-3) / 1e12, "frac": tot_flop / (ms * 1e-3) / 1e12 / peak,
                          "fp64_pipe_busy_time_weighted": tot_busy / ms, "kernel_ms": tot_ms / args.steps, "ms": ms / args.steps}
        rf["reference_equivalent_x"] = (f_tab * n * args.steps) / (tm["ms_alpha"] * 1e-3) / 1e12 / peak if tm["ms_alpha"] > 0 else None
        line = {
            "metric": "(mphi,g) evolutions/sec at N_E=200", "value": value, "unit": "evolutions/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(world),
            "us_per_z_step": 1e3 * tm["ms_integrate"] / (n * args.steps * (Nz - 1)),
            "kernel_ms_per_step": {k: v / args.steps for k, v in ms_kernel.items()},
            "roofline": rf,
            "e2e": {"value": n * world * e2e_steps / e2e_s, "unit": "evolutions/s", "h2d_bytes_per_step": int(3 * 8 * n),
                    "d2h_bytes_per_step": int(2 * 3 * N * 8 * n + 4 * n), "bitwise_equal_to_device_path": same},
            "gpu_launches": tm["n_launches"],
            "outputs_checked": check,
            "clocks": clocks,
        }
        if world == 1 and not args.points:
            # parity of THIS workload, in this run: the whole scan shard by shard in scan order (default builder), against
            # the oracle fixture of the seeded 1024-point subsample -- outside the timed region
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            try:
                import scan_parity as SP
                got, st, info = SP.evolve_fixture_points_in_their_shards(ctx)
                rep = SP.evaluate(got)
                rep["grouping_per_shard"] = info
                rep["fixture"] = "tests/golden/oracle_scan1024.npz (strict oracle + 6-member ensemble; tools/gen_scan_golden.py)"
                line["parity"] = rep
            except Exception as ex:
                line["parity"] = {"unavailable": str(ex)}
        if world == 1 and not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            cnt = max(cores * 32, 32)
            try:
                wall = run_reference_sample(cpu_sample_points(cnt), cores)
                line["cpu_baseline"] = {"value": cnt / wall, "unit": "evolutions/s", "cores": cores, "kind": "reference",
                                        "sample": "%d evolve() of the seeded scan subsample, one process per point, %d concurrent, %.1f s" % (cnt, cores, wall)}
            except Exception as ex:  # oracle binary absent
                line["cpu_baseline"] = {"value": None, "unit": "evolutions/s", "cores": cores, "kind": "reference", "sample": "unavailable: %s" % ex}
        print(json.dumps(line))
    run.close()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


# ---------------------------------------------------------------------------------------------
# CUDA arm, BASELINE configs[3]: fine binning (N_E = 2000).  A side line, not the headline metric.
# ---------------------------------------------------------------------------------------------
def bench_fine(args):
    import torch
    import nusiprop_b200 as nb

    local = int(os.environ.get("LOCAL_RANK", "0"))
    if int(os.environ.get("RANK", "0")) != 0:
        return 0
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    ctx = nb.Context(majorana=True, non_resonant=True, normal_ordering=True, N_bins_E=FINE_N_E, lEmin=LEMIN, lEmax=LEMAX, zmax=ZMAX,
                     flav=FLAV, phiphi=False, device=local)
    N, Nz, NT = ctx.N_bins_E, ctx.N_steps_z, ctx.NT
    npts = args.points or 148
    # a small scan: rows of 8 couplings per mediator mass; the first point is BASELINE's (100, 1e-2)
    nm = (npts + 7) // 8
    mphi = np.repeat(100.0 * 10 ** (np.arange(nm) * 0.05), 8)[:npts]
    g = np.tile(1e-2 * 10 ** (np.arange(8) * 0.25), nm)[:npts]
    mntot = np.full(npts, MNTOT)
    pts = ctx.upload_points(mphi, g, mntot)
    fm = torch.empty((npts, 3, N), dtype=torch.float64, device=dev)
    ff = torch.empty_like(fm)
    st = torch.zeros(npts, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    steps, warm = max(1, min(args.steps, 3)), max(1, min(args.warmup, 3))
    for _ in range(warm):
        ctx.evolve_points(pts, fm.data_ptr(), ff.data_ptr(), st.data_ptr(), stream)
    torch.cuda.synchronize()
    ctx.last_timing()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        ctx.evolve_points(pts, fm.data_ptr(), ff.data_ptr(), st.data_ptr(), stream)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    tm = ctx.last_timing()
    clocks = sampler.stop()
    check = validate_outputs(fm.cpu().numpy(), ff.cpu().numpy(), st.cpu().numpy(), "fine workload")
    # one point alone: the latency of a single N_E = 2000 evolve() (what a drop-in calculate_flux::evolve() costs)
    t0 = time.perf_counter()
    one = ctx.evolve(100.0, 1e-2, MNTOT)
    t_first = time.perf_counter() - t0
    ctx.last_timing()  # the kernel times below are those of the second call alone
    t0 = time.perf_counter()
    one = ctx.evolve(100.0, 1e-2, MNTOT)
    t_one = time.perf_counter() - t0
    tm1 = ctx.last_timing()
    table_bytes = NT * (NT - 1) / 2 * 8
    hbm = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs") if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else None
    n_bands = (Nz - 1 + 31) // 32
    int_gbs = table_bytes * n_bands * npts * steps / (tm["ms_integrate"] * 1e-3) / 1e9 if tm["ms_integrate"] > 0 else None
    int_flop = 2.0 * N * (N - 1) / 2 * (Nz - 1)
    int_tf = int_flop * npts * steps / (tm["ms_integrate"] * 1e-3) / 1e12 if tm["ms_integrate"] > 0 else None
    peak = nb.measure_fp64_peak(local)
    line = {
        "metric": "(mphi,g) evolutions/sec at N_E=2000", "value": npts * steps / (ms * 1e-3), "unit": "evolutions/s", "n_gpus": 1,
        "steps": steps, "warmup": warm, "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": fine_config(npts),
        "us_per_z_step": 1e3 * tm["ms_integrate"] / (npts * steps * (Nz - 1)),
        "kernel_ms_per_step": {"rows_kernel": tm["ms_rows"] / steps, "alpha_kernel": tm["ms_alpha"] / steps, "integrate_kernel": tm["ms_integrate"] / steps},
        "single_point": {"seconds_wall": t_one, "seconds_wall_first_call": t_first, "integrate_ms": tm1["ms_integrate"], "alpha_ms": tm1["ms_alpha"], "rows_ms": tm1["ms_rows"],
                         "us_per_z_step": 1e3 * tm1["ms_integrate"] / (Nz - 1), "finite": bool(np.isfinite(one[0]).all())},
        # the integrator dominates this configuration.  Its arithmetic is one FMA per (target bin, source bin, z-step):
        # N(N-1)/2 x (N_z-1) FMAs per point on the FP64 pipe; its memory stream is the packed table, once per band of 32
        # z-steps.  Neither is the bound: the wave-to-wave dependency chain is (N + 31 waves per band, one warp).
        "roofline": {"bound": "fp64", "kernel": "integrate_cluster_kernel", "achieved": int_tf, "peak": peak, "unit": "TFLOP/s",
                     "frac": (int_tf / peak) if (int_tf and peak) else None, "traffic": None,
                     "peak_source": "DFMA rate measured in this run by nusi_measure_fp64_peak",
                     "note": "algorithmic flops = 2 x N(N-1)/2 x (N_z-1) = %.3f GFLOP per point / the integrator's CUDA-event time" % (int_flop / 1e9),
                     "hbm": {"achieved": int_gbs, "peak": hbm, "unit": "GB/s", "frac": (int_gbs / hbm) if (int_gbs and hbm) else None,
                             "note": "packed alpha table (%.1f MB per point) read once per band of 32 z-steps (%d bands) / the same time; peak = MEASURED_PEAKS.json hbm_gbs" % (table_bytes / 1e6, n_bands)}},
        "gpu_launches": tm["n_launches"], "outputs_checked": check, "clocks": clocks,
    }
    print(json.dumps(line))
    pts.close()
    ctx.close()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--workload", default="scan", choices=["scan", "fine"], help="scan = the headline 256x256 scan at N_E=200; fine = N_E=2000 (configs[3])")
    ap.add_argument("--points", type=int, default=0, help="debug: evolve only the first POINTS points of the shard (fine: batch size, default 148)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return bench_reference(args)
    if args.workload == "fine":
        return bench_fine(args)
    return bench_cuda(args)


if __name__ == "__main__":
    sys.exit(main())
