#!/usr/bin/env python
"""bench.py — RHS evaluations/s and rays/s of the batched whistler ray tracer on N B200s.

Workload (config.workload): "cfg3" (default) = BASELINE.json configs[2], the 3D (RayTrace_3D.jl) sweep of 10^6 rays
over colatitude x longitude x wave-normal angle that north_star's target is quoted on.  At N GPUs the sweep is refined
along longitude to N*10^6 rays and partitioned across ranks (rotated-strided; weak scaling, no data-path collective);
the one exchange is the gather of terminal records to rank 0 (inside the timed step for N > 1).
Other workloads (records under profiles/, not the driver's line): "cfg2" = configs[1], 10^5-ray 2D sweep; "cfg4" =
configs[3], 10^6 long-duration 2D rays (tspan 5e10, reltol 1e-7, maxiters 10^6: minutes per step, run it with
--steps 1 --warmup 0); "cfg5" = configs[4], the FIXED 10^7-ray 3D list partitioned over the ranks (strong scaling).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg3|cfg2|cfg4|cfg5] [--impl ours|reference]

A step = one full solve of the rank's launch list (init kernel + integrate slices + group-delay kernel).
value      : RHS evaluations/s, launch list and results resident in HBM (device-pointer C-ABI entry point)
e2e        : same metric through rt_solve_3d with pinned HOST buffers (H2D + D2H inside the timed region)
roofline   : FP64 CUDA-core FLOP/s of the integrate kernel vs the DFMA peak measured in this run
cpu_baseline / --impl reference : the CPU oracle (a port of the Julia path; Julia is not available) on all host cores
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

RE = 6.3712e6
# Algorithmic FP64 work per Haselgrove RHS evaluation: executed
# DADD + DMUL + 2*DFMA per RHS of the minimal faithful schedule, calibrated once with ncu
# (profiles/r1_flop_calibration.md) and frozen.
FLOP_PER_RHS = {"cfg3": 2602.0, "cfg5": 2602.0, "cfg2": 2307.0, "cfg4": 2307.0, "cfg2save": 2307.0}
IS3D = {"cfg3": True, "cfg5": True, "cfg2": False, "cfg4": False, "cfg2save": False}
SCALING = {"cfg3": "weak", "cfg2": "weak", "cfg4": "weak", "cfg5": "strong", "cfg2save": "weak"}
# "cfg2save": the cfg2 launch list with the reference's save-every-step output on (RayTrace_lat.jl:318-329,345-372):
# 8192 rows of 72 B per ray, iteration cap 8190 so that no trajectory is truncated -> 59 GB of trajectory rows per step
SAVE_ROWS = 8192
# dram__bytes_read.sum + dram__bytes_write.sum of ONE full-list launch of rt_integrate_kernel (4 attempts for each of
# 10^6 rays, 7813 blocks) from the ncu --set full capture summarised in profiles/r2_ncu_integrate.md:
# 176.4 MB + 120.4 MB.  Algorithmic traffic of that launch: ray state in and out, 2 x ~200 B x 10^6 = 0.4 GB upper bound.
# launch-list partition across ranks: block b of N consecutive rays goes to the ranks rotated by b, so that every GPU
# gets its share of the long rays AND of every wave-normal angle (plain strided: +-4.5 % work per rank on 8 GPUs)
PARTITION = "rotated"


def partition_layout(world, is3d):
    """Across ranks: blocks of 4096 consecutive rays go round the ranks (as rt_solve_*_multi does inside the library).
    Work per rank within 0.5 % on the 8 x 10^6-ray sweep, and the root places every rank's records with block copies:
    measured on 8 GPUs 55.2 ms per step (gather: 1.0 ms receive + 0.4 ms placement) against 56.0 ms with the per-ray
    rotated layout (1.2 ms of indexed placement) — profiles/r2_bench_cfg3_8gpu_layout_*.json."""
    if LAYOUT_OVERRIDE:
        return LAYOUT_OVERRIDE
    return "blockcyclic:4096" if world > 1 else PARTITION
NCU_DRAM_BYTES_PER_FULL_LAUNCH = {"cfg3": 296.8e6, "cfg5": None, "cfg2": 13.3e6, "cfg4": None, "cfg2save": None}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="cfg3", choices=["cfg3", "cfg2", "cfg4", "cfg5", "cfg2save"])
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rays-per-gpu", type=int, default=0, help="override (testing only)")
    ap.add_argument("--slice", type=int, default=0)
    ap.add_argument("--coop-threshold", type=int, default=0, help="developer: rt_solver_opts.coop_threshold (0 = default)")
    ap.add_argument("--layout", default="", help="developer: launch-list partition across ranks (default: see partition_layout)")
    ap.add_argument("--cpu-sample", type=int, default=0, help="rays in the CPU sample (0 = default)")
    ap.add_argument("--maxiters", type=int, default=0, help="iteration cap override (a cut-off run: named in config)")
    ap.add_argument("--skip-e2e", action="store_true",
                    help="do not repeat the steps through the host-pointer entry point (cfg4: a step is minutes long); e2e is null")
    return ap.parse_args()


def make_launch(workload, world, rank, rays_per_gpu=0):
    """This rank's partition of the (weak-scaled) sweep; returns (launch [ncol, n_local], n_total)."""
    from raytrace_b200 import sweeps
    if workload == "cfg3":
        if rays_per_gpu:
            k = max(1, int(round((rays_per_gpu / 100.0) ** 0.5)))
            full = sweeps.sweep_3d(k, k * world, 100)
        else:
            full = sweeps.sweep_3d(100, 100 * world, 100)
    elif workload == "cfg5":  # the fixed 10^7-ray list (cfg3's grid refined along longitude), whatever the world size
        full = sweeps.sweep_3d(100, max(1, rays_per_gpu // 10000) if rays_per_gpu else 1000, 100)
    elif workload == "cfg4":
        if rays_per_gpu:
            full = sweeps.sweep_2d_long(max(1, rays_per_gpu // 1000) * world, 1000)
        else:
            full = sweeps.sweep_2d_long(1000 * world, 1000)
    else:
        if rays_per_gpu:
            full = sweeps.sweep_2d(max(1, rays_per_gpu // 250) * world, 250)
        else:
            full = sweeps.sweep_2d(400 * world, 250)
    n_total = full.shape[1]
    layout = partition_layout(world, IS3D[workload])
    return np.ascontiguousarray(full[:, sweeps.partition_indices(n_total, rank, world, layout)]), n_total


MAXITERS_OVERRIDE = 0
COOP_THRESHOLD = 0
LAYOUT_OVERRIDE = ""


def solver_kwargs(workload, slice_attempts=0):
    """Solver options of the named config."""
    if IS3D[workload]:  # RayTrace_3D.jl:395 defaults
        kw = dict(adaptive=1, reltol=1e-3, abstol=1e-6, dtmax=0.0, dtmin=0.0, maxiters=100000, t1=5.0e9, r_esc=10 * RE)
    elif workload == "cfg4":  # scratch.jl:390 tspan, RayTrace_lat.jl:338 tolerances, maxiters 10^6
        kw = dict(adaptive=1, reltol=1e-7, abstol=1e-6, dtmax=1e6, dtmin=1e-8, maxiters=1000000, t1=5.0e10, r_esc=10 * RE)
    else:  # RayTrace_lat.jl:338
        kw = dict(adaptive=1, reltol=1e-7, abstol=1e-6, dtmax=1e6, dtmin=1e-8, maxiters=100000, t1=5.0e9, r_esc=10 * RE)
    if workload == "cfg2save":
        kw.update(maxiters=SAVE_ROWS - 2, save_mode=1, max_samples=SAVE_ROWS, save_stride=1)
    if MAXITERS_OVERRIDE:
        kw["maxiters"] = MAXITERS_OVERRIDE
    if slice_attempts:
        kw["slice_attempts"] = slice_attempts
    if COOP_THRESHOLD:
        kw["coop_threshold"] = COOP_THRESHOLD
    return kw


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1]))
                mx.append(float(p[2]))
            except ValueError:
                continue
            for nm, v in zip(names, p[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_oracle_run(workload, launch, nthreads):
    """Time the CPU oracle (port of the Julia path) on `launch`; returns (seconds, total RHS evals, n rays)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    kw = solver_kwargs(workload)
    kw.pop("adaptive")
    kw.pop("coop_threshold", None)  # scheduling only
    o = O.opts(**kw)
    E = O.env(O.RAY3D if IS3D[workload] else O.LAT2D)
    t = time.perf_counter()
    out = O.solve_batch(E, o, launch, IS3D[workload], nthreads)
    dt = time.perf_counter() - t
    return dt, int(out["nfe"].astype(np.int64).sum()), launch.shape[1]


def cpu_sample(workload, launch, nrays):
    """A bounded sample of the rank-0 launch list: ray k * 49999 mod n (a stride coprime to the sweeps' grid sizes, so
    that every wave-normal angle, longitude and latitude is drawn — an even stride lands on 2 of cfg3's 100 angles)."""
    n = launch.shape[1]
    if nrays >= n:
        return launch
    stride = 49999
    while np.gcd(stride, n) != 1:
        stride += 2
    idx = (np.arange(nrays, dtype=np.int64) * stride) % n
    return np.ascontiguousarray(launch[:, idx])


CPU_SAMPLE_DEFAULT = {"cfg3": 40000, "cfg5": 40000, "cfg2": 96, "cfg4": 32, "cfg2save": 256}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path (oracle port; Julia is unavailable on this image) on all host
    cores, each step a bounded sample of the same workload.  Rank 0 only."""
    if rank != 0:
        return
    launch, _ = make_launch(args.workload, 1, 0, args.rays_per_gpu)
    ncores = os.cpu_count() or 1
    nsample = args.cpu_sample or CPU_SAMPLE_DEFAULT[args.workload]
    sample = cpu_sample(args.workload, launch, nsample)
    for _ in range(min(args.warmup, 1)):
        cpu_oracle_run(args.workload, cpu_sample(args.workload, launch, max(64, nsample // 20)), ncores)
    tot_t, tot_nfe, tot_rays = 0.0, 0, 0
    for _ in range(args.steps):
        dt, nfe, nr = cpu_oracle_run(args.workload, sample, ncores)
        tot_t += dt
        tot_nfe += nfe
        tot_rays += nr
    val = tot_nfe / tot_t
    desc = f"{sample.shape[1]} rays (index k * 49999 mod n) of the {launch.shape[1]}-ray {args.workload} launch list per step"
    line = {
        "impl": "reference", "metric": "rhs_evals_per_s", "value": val, "unit": "RHS evals/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / args.steps, "higher_is_better": True,
        "scaling": SCALING[args.workload], "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "rays_per_s": tot_rays / tot_t,
        "config": {"workload": args.workload, "rays_per_step": int(sample.shape[1]), "solver": solver_kwargs(args.workload)},
        "cpu_baseline": {"value": val, "unit": "RHS evals/s", "cores": ncores, "kind": "port", "sample": desc},
        "e2e": {"value": val, "unit": "RHS evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "reference is Julia (no julia binary here); this is the plain-C oracle port, which calls the "
                "refractive index 12x (3D) / 9x (2D) per RHS exactly as the .jl files do",
    }
    print(json.dumps(line), flush=True)


def main():
    global MAXITERS_OVERRIDE, COOP_THRESHOLD, LAYOUT_OVERRIDE
    args = parse()
    LAYOUT_OVERRIDE = args.layout
    MAXITERS_OVERRIDE = args.maxiters
    COOP_THRESHOLD = args.coop_threshold
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    import raytrace_b200 as R
    from raytrace_b200 import multi

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    launch_h, n_total = make_launch(args.workload, world, rank, args.rays_per_gpu)
    ncol, n = launch_h.shape
    is3d = IS3D[args.workload]
    dim = 7 if is3d else 4
    rt = R.RayTracer("ray3d" if is3d else "lat2d", device=local_rank)
    rt.set_solver(**solver_kwargs(args.workload, args.slice))
    stream = torch.cuda.current_stream()
    rt.set_stream(stream.cuda_stream)
    peak = rt.measure_fp64_peak()

    # device-resident launch list and results
    launch_d = torch.from_numpy(launch_h).to(dev)
    # the solver writes straight into the two gather buffers (doubles [dim+2, n], int32 [5, n]): no packing step
    rec_d, ints, views = multi.alloc_terminal(n, dim, dev)
    u_end, t_end, gd = views["u_end"], views["t_end"], views["group_delay"]
    res_ptrs = {k: v.data_ptr() for k, v in views.items()}
    cols = [launch_d[k].data_ptr() for k in range(ncol)]
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def step_device():
        st = rt.solve_device(n, cols, res_ptrs)
        if world > 1:
            multi.gather_terminal(rec_d, ints, n_total, rank, world, layout=partition_layout(world, is3d))
        return st

    for _ in range(args.warmup):
        step_device()
    peak = max(peak, rt.measure_fp64_peak())  # again with the clocks up: the larger figure is the denominator
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    launches, integ_ms, slices = 0, 0.0, 0
    for k in range(args.steps):
        flush.fill_(k & 0xFF)  # L2 flush between timed iterations (not timed)
        barrier()
        ev[k][0].record(stream)
        st = step_device()
        ev[k][1].record(stream)
        launches += st.kernel_launches
        integ_ms += st.integrate_ms
        slices += st.slices
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    peak = max(peak, rt.measure_fp64_peak())
    ms_local = sum(a.elapsed_time(b) for a, b in ev)
    nfe_local = int(ints[3].to(torch.int64).sum().item())
    status_h = ints[0].cpu().numpy()
    # the device-resident step's records, kept for the comparison with the host-pointer path below
    dev_res = {"u_end": u_end.cpu().numpy(), "t_end": t_end.cpu().numpy(), "group_delay": gd.cpu().numpy(),
               "status": status_h, "naccept": ints[1].cpu().numpy(), "nreject": ints[2].cpu().numpy(),
               "nfe": ints[3].cpu().numpy(), "nreflect": ints[4].cpu().numpy()}
    # diagnostic, outside the timed region: what the gather costs its root (all ranks take part)
    gather_ms = None
    if world > 1:
        gather_ms = {}
        for _ in range(3):
            barrier()
            multi.gather_terminal(rec_d, ints, n_total, rank, world, layout=partition_layout(world, is3d), timing=gather_ms)

    # trajectory output (save workloads): rows written per step, the same sweep without saving, and the CSR packing
    traj = None
    if rank == 0 and solver_kwargs(args.workload).get("save_mode"):
        lib = R.api._lib()
        total = ctypes.c_int64()
        if lib.rt_traj_size(rt._ctx, ctypes.byref(total)) != 0:
            raise RuntimeError(lib.rt_last_error(rt._ctx).decode())
        ncol_t = lib.rt_traj_columns(rt._ctx)
        tbytes = total.value * ncol_t * 8
        flags = np.zeros(n, np.int32)
        lib.rt_traj_truncated(rt._ctx, flags.ctypes.data_as(ctypes.c_void_p))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        d_off, d_rows, tot2 = ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_int64()
        pack_ms = []
        for _ in range(3):
            e0.record(stream)
            if lib.rt_traj_pack_device(rt._ctx, ctypes.byref(d_off), ctypes.byref(d_rows), ctypes.byref(tot2)) != 0:
                raise RuntimeError(lib.rt_last_error(rt._ctx).decode())
            e1.record(stream)
            torch.cuda.synchronize()
            pack_ms.append(e0.elapsed_time(e1))
        kw0 = dict(solver_kwargs(args.workload, args.slice), save_mode=0, max_samples=0)
        rt.set_solver(**kw0)
        rt.solve_device(n, cols, res_ptrs)
        st0 = rt.solve_device(n, cols, res_ptrs)
        rt.set_solver(**solver_kwargs(args.workload, args.slice))
        step_ms = ms_local / args.steps
        traj = {"rows_per_step": int(total.value), "bytes_per_step": int(tbytes), "row_bytes": ncol_t * 8,
                "truncated_rays": int(flags.sum()),
                "write_gbs_over_the_step": tbytes / (step_ms * 1e-3) / 1e9,
                "step_ms_saving": step_ms, "step_ms_not_saving": st0.total_ms,
                "pack_ms": min(pack_ms), "pack_gbs": 2 * tbytes / (min(pack_ms) * 1e-3) / 1e9,
                "pack_note": "rt_traj_pack_device: cub scan of the row counts + one coalesced copy kernel; "
                             "bytes read + written over its time, against hbm_gbs of MEASURED_PEAKS.json",
                "hbm_gbs_measured": json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs")
                if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else None}

    # diagnostic, outside the timed region: the integrate kernel in steady state (every ray alive, full lists):
    # the same sweep cut off after 24 step attempts, one chain
    steady = None
    if rank == 0:
        kw = solver_kwargs(args.workload, args.slice)
        rt.set_solver(**dict(kw, maxiters=24, single_chain=1))
        rt.solve_device(n, cols, res_ptrs)
        st2 = rt.solve_device(n, cols, res_ptrs)
        nfe2 = int(ints[3].to(torch.int64).sum().item())
        steady = {"rhs_evals_per_s": nfe2 / (st2.integrate_ms * 1e-3), "launches": int(st2.slices),
                  "how": "same launch list, maxiters=24, single chain: integrate launches over full work lists only"}
        rt.set_solver(**dict(kw, single_chain=0))

    # e2e: pinned host buffers through the host-pointer entry point (H2D + D2H inside the timed region)
    launch_p = torch.from_numpy(launch_h).pin_memory()
    res_h = R.api.SolveResult(n, dim)
    pinned = {}
    for name in ("u_end", "t_end", "group_delay", "status", "naccept", "nreject", "nfe", "nreflect"):
        t = torch.from_numpy(getattr(res_h, name)).pin_memory()
        pinned[name] = t
        setattr(res_h, name, t.numpy())
    cres = res_h._c()
    fn = R.api._lib().rt_solve_3d if is3d else R.api._lib().rt_solve_2d
    hcols = [ctypes.c_void_p(launch_p[k].data_ptr()) for k in range(ncol)]

    def step_host():
        rc = fn(rt._ctx, ctypes.c_int64(n), *hcols, ctypes.byref(cres))
        if rc != 0:
            raise RuntimeError(R.api._lib().rt_last_error(rt._ctx).decode())
        if world > 1:  # host-side assembly on rank 0 is the caller's job; keep ranks in step
            dist.barrier()

    if not args.skip_e2e:
        step_host()
    barrier()
    e2e_ms = 0.0
    for k in range(0 if args.skip_e2e else args.steps):
        flush.fill_(k & 0xFF)
        barrier()
        t0 = time.perf_counter()
        step_host()
        torch.cuda.synchronize()
        e2e_ms += 1e3 * (time.perf_counter() - t0)
    barrier()
    # the host-pointer path (early result copy + late patch) must deliver what the device-resident path delivered
    for name, b in ({} if args.skip_e2e else dev_res).items():  # every record, bit for bit (NaN == NaN)
        a = getattr(res_h, name)
        if not np.array_equal(a.view(np.uint64) if a.dtype == np.float64 else a,
                              b.view(np.uint64) if b.dtype == np.float64 else b):
            raise RuntimeError(f"e2e results differ from the device-resident solve in {name}")
    h2d = ncol * n * 8
    d2h = (dim + 2) * n * 8 + 5 * n * 4

    # max over ranks / sums over ranks
    if world > 1:
        tmax = torch.tensor([ms_local, e2e_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = torch.tensor([nfe_local, n, launches], dtype=torch.float64, device=dev)
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        ms_all, e2e_all = float(tmax[0]), float(tmax[1])
        nfe_all, rays_all, launches_all = float(tsum[0]), float(tsum[1]), float(tsum[2])
        per_rank = [torch.zeros(2, dtype=torch.float64, device=dev) for _ in range(world)]
        dist.all_gather(per_rank, torch.tensor([ms_local, nfe_local], dtype=torch.float64, device=dev))
        by_rank = [{"ms_per_step": float(t[0]) / args.steps, "rhs_evals_per_step": int(t[1])} for t in per_rank]
    else:
        ms_all, e2e_all, nfe_all, rays_all, launches_all = ms_local, e2e_ms, float(nfe_local), float(n), float(launches)
        by_rank = None

    if rank == 0:
        K = args.steps
        value = nfe_all * K / (ms_all * 1e-3)
        flop = FLOP_PER_RHS[args.workload]
        # dominant kernel = rt_integrate_kernel (> 99 % of the step): algorithmic FLOP of rank 0's step over rank 0's
        # WHOLE step time (init kernel, every integrate slice, group-delay kernel, scheduling gaps)
        achieved = (nfe_local * K * flop) / (ms_local * 1e-3) / 1e12
        ncores = os.cpu_count() or 1
        cpu = None
        if world == 1 or True:
            nsample = args.cpu_sample or CPU_SAMPLE_DEFAULT[args.workload]
            sample = cpu_sample(args.workload, launch_h, nsample)
            cdt, cnfe, cn = cpu_oracle_run(args.workload, sample, ncores)
            cpu = {"value": cnfe / cdt, "unit": "RHS evals/s", "cores": ncores, "kind": "port",
                   "sample": f"{cn} rays strided from rank 0's {n}-ray launch list, {cdt:.1f} s wall",
                   "rays_per_s": cn / cdt}
        uniq, cnt = np.unique(status_h, return_counts=True)
        line = {
            "metric": "rhs_evals_per_s", "value": value, "unit": "RHS evals/s", "n_gpus": world, "steps": K,
            "warmup": args.warmup, "ms_per_step": ms_all / K, "higher_is_better": True, "scaling": SCALING[args.workload],
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "rays_per_s": rays_all * K / (ms_all * 1e-3),
            "config": {"workload": args.workload, "rays_total": int(rays_all), "rays_per_gpu": int(n),
                       "rhs_evals_per_step": int(nfe_all), "solver": solver_kwargs(args.workload, args.slice),
                       "cut_off": bool(args.maxiters), "partition": f"{partition_layout(world, is3d)} x{world}", "l2": "flushed between timed steps (256 MiB fill)",
                       "status_counts_rank0": {R.STATUS_NAMES[int(a)]: int(b) for a, b in zip(uniq, cnt)}},
            "roofline": {"bound": "fp64", "achieved": achieved, "peak": peak / 1e12, "unit": "TFLOP/s",
                         "frac": achieved / (peak / 1e12), "traffic": NCU_DRAM_BYTES_PER_FULL_LAUNCH[args.workload],
                         "traffic_note": "bytes per full-list launch of the 10^6-ray list (ncu --set full, "
                                         "profiles/r1_ncu_integrate_v3.md); 0.09 TB/s: HBM is not the bound",
                         "steady_state": None if steady is None else dict(
                             steady, achieved=steady["rhs_evals_per_s"] * flop / 1e12,
                             frac=steady["rhs_evals_per_s"] * flop / peak),
                         "peak_source": "DFMA-chain microbenchmark in this run (MEASURED_PEAKS.json has no FP64 entry)",
                         "flop_per_rhs": flop, "kernel": "rt_integrate_kernel",
                         "how": "rhs_evals_per_step x flop_per_rhs / ms_per_step of rank 0 (whole step)",
                         "integrate_span_ms_per_step": integ_ms / K, "launches": int(slices)},
            "cpu_baseline": cpu,
            "e2e": None if args.skip_e2e else {
                "value": nfe_all * K / (e2e_all * 1e-3), "unit": "RHS evals/s", "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": int(d2h), "ms_per_step": e2e_all / K, "rays_per_s": rays_all * K / (e2e_all * 1e-3)},
            "trajectory": traj,
            "gather_root_ms": gather_ms,
            "gpu_launches": int(launches_all),
            "by_rank": by_rank,
            "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    rt.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
