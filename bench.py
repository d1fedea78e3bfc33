#!/usr/bin/env python
"""Benchmark of the FV hot path (BASELINE.json metric: FV cell-updates/s, % HBM roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload at N=1: BASELINE config 2, the 2D parabolic slider bearing on 4096 x 4096 cells,
FP64, Dowson-Higginson EOS, fixed-form stresses, x: Dirichlet(rho)/Neumann(jx, jy), y
periodic, adaptive time step.  N>1: the same slider slab-decomposed
along x with 4096 x 4096 cells per GPU (weak scaling).

One "step" = one Problem.update(): predictor pass + corrector pass (+ temporal average,
validity test, reductions) + the scalar bookkeeping kernel.  `value` times K steps with the
fields resident in HBM; `e2e` times the host-buffer entry point (pinned host q uploaded
and downloaded every step).  The oracle is used only for the cpu_baseline / --impl
reference legs.
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

N_CELLS_1D = 4096            # cells per GPU per axis
BYTES_PER_CELL_UPDATE = 168  # 21 doubles (pass 1: 9, pass 2: 12)
BYTES_PRED, BYTES_CORR = 72, 96

SLIDER_YAML = """
options:
    output: data/bench
    write_freq: 1000000
    silent: True
grid:
    Lx: {Lx}
    Ly: 0.0762
    Nx: {Nx}
    Ny: {Ny}
    xE: ['D', 'N', 'N']
    xW: ['D', 'N', 'N']
    yS: ['P', 'P', 'P']
    yN: ['P', 'P', 'P']
    xE_D: 877.7007
    xW_D: 877.7007
geometry:
    type: parabolic
    hmin: 2.54e-5
    hmax: 5.08e-5
    U: 4.57
    V: 0.
numerics:
    CFL: 0.45
    adaptive: 1
    tol: 1e-12
    dt: 1e-10
    max_it: 100000000
properties:
    shear: 0.0794
    bulk: 0.
    EOS: DH
    P0: 101325.
    rho0: 877.7007
    C1: 3.5e10
    C2: 1.23
"""


def slider_inputs(Nx, Ny):
    from gapflow_b200 import config
    return config.read_yaml_string(SLIDER_YAML.format(Nx=Nx, Ny=Ny, Lx=0.0762 * Nx / Ny))


def initial_fields(d, seed=1234, rows=None):
    """_initialize + the seeded perturbation + topography, host NumPy.

    ``rows=(a, b)``: padded x-rows a..b-1 of the global field only (one rank's slab).  The
    perturbation is drawn per global row from its own generator, so every decomposition sees
    the same global field; ghost cells are filled by the solver afterwards."""
    import numpy as np
    from gapflow_b200.topography import make_topography
    g, geo, prop = d['grid'], d['geometry'], d['properties']
    a, b = rows if rows is not None else (0, g['Nx'] + 2)
    shape = (b - a, g['Ny'] + 2)
    q = np.empty((3, *shape))
    q[0] = prop['rho0']
    q[1] = prop['rho0'] * geo['U'] / 2.0
    q[2] = prop['rho0'] * geo['V'] / 2.0
    blk = 256                                           # rows per generator
    for b0 in range(a // blk * blk, b, blk):
        rng = np.random.default_rng([seed, b0 // blk])
        n = rng.standard_normal((3, blk, shape[1]))
        lo, hi = max(a, b0), min(b, b0 + blk)
        sl, ns = slice(lo - a, hi - a), slice(lo - b0, hi - b0)
        q[0, sl] *= 1 + 1e-6 * n[0, ns]
        q[1, sl] += 1e-3 * prop['rho0'] * geo['U'] * n[1, ns]
        q[2, sl] += 1e-3 * prop['rho0'] * geo['U'] * n[2, ns]
    return q, make_topography(g, geo, rows=rows)


class ClockSampler:
    """Samples SM clocks and throttle reasons with nvidia-smi while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            self.t.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except (ValueError, IndexError):
                pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------
# CPU legs (the only place the oracle is executed by this file)
# ------------------------------------------------------------------------------------------
def cpu_steps(n_side, steps, warmup):
    """Times the NumPy oracle (port of the reference's update loop, 1 core: NumPy ufuncs are
    single-threaded and so is the reference) on an n_side^2 sample of the slider workload."""
    import numpy as np
    from oracle.fv import FVOracle
    d = slider_inputs(n_side, n_side)
    q, topo = initial_fields(d)
    o = FVOracle(d['grid'], d['geometry'], d['properties'], d['numerics'], topo=topo)
    o.q[...] = q
    o.fill_ghosts()
    o.dt = d['numerics']['CFL'] * o.dt_crit
    for _ in range(warmup):
        o.update()
    t0 = time.perf_counter()
    for _ in range(steps):
        o.update()
    dt = time.perf_counter() - t0
    assert np.isfinite(o.kinetic_energy)
    return n_side * n_side * steps / dt, dt / steps


def pick_cpu_sample(budget_s, steps):
    """Largest power-of-two grid whose `steps` oracle steps fit the budget (calibrated on 256^2)."""
    rate, _ = cpu_steps(256, 1, 1)
    n = 256
    while n < 2048 and (2 * n) ** 2 * steps / (0.6 * rate) < budget_s:   # big grids run ~40 % slower per cell
        n *= 2
    return n


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (NumPy, host cores)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = args.steps, args.warmup
    n = pick_cpu_sample(150.0, steps + warmup)
    rate, s_per_step = cpu_steps(n, steps, warmup)
    line = {
        "impl": "reference", "metric": "FV cell-updates/s", "value": rate, "unit": "cell-updates/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": 1e3 * s_per_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "2D parabolic slider bearing, DH EOS, fixed-form stress, FP64 "
                               f"(BASELINE config 2: 4096x4096 per GPU); CPU sample {n}x{n} of the same problem",
                   "l2": "n/a (CPU)"},
        "cpu_baseline": {"value": rate, "unit": "cell-updates/s", "cores": 1, "kind": "port",
                         "sample": f"{n}x{n} cells, {steps} steps after {warmup} warm-up; NumPy restatement of "
                                   "the reference's update loop (its ufuncs are single-threaded); "
                                   f"host has {os.cpu_count()} logical cores"},
        "e2e": {"value": rate, "unit": "cell-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------
def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def run_gpu(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from gapflow_b200 import _capi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = _capi.load()

    n = N_CELLS_1D
    d = slider_inputs(n * world, n)
    if world == 1:
        from gapflow_b200.fv import FVSolver
        q0, topo = initial_fields(d)
        solver = FVSolver(d['grid'], d['geometry'], d['properties'], d['numerics'], topo=topo, arith=args.arith)
        solver.set_q(q0)
        solver.fill_ghosts()
        solver.pre_run(False)
        step = solver.step
    else:
        from gapflow_b200.dist import SlabSolver, slab_rows
        start, nloc = slab_rows(d['grid']['Nx'], world)[rank]
        q0, topo = initial_fields(d, rows=(start, start + nloc + 2))
        solver = SlabSolver(d['grid'], d['geometry'], d['properties'], d['numerics'], topo=topo, q=q0,
                            rank=rank, world=world, arith=args.arith,
                            overlap=os.environ.get('GPF_SLAB_OVERLAP', '1') != '0')
        solver.fill_ghosts()
        solver.pre_run(False)
        step = solver.step
        q0 = None
    cells_per_step = n * n * world

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput --------------------------------------------------------
    step(args.warmup)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = lib.gpf_launch_count()
    with ClockSampler(local) as clk:
        barrier()
        e0.record()
        t_host = time.perf_counter()
        step(args.steps)
        t_host = time.perf_counter() - t_host          # enqueue only: the host does not wait for the GPU
        e1.record()
        barrier()
    launches = lib.gpf_launch_count() - l0
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    sc = solver.scalars()
    assert sc.step == args.warmup + args.steps and not sc.stop, "benchmark run went invalid"
    value = cells_per_step * args.steps / (ms * 1e-3)

    # ---- per-kernel durations (CUDA events on the launching stream, separate batch) --------
    ks = min(args.steps, 20)
    solver.set_timing(True)
    step(ks)
    t_pred, t_corr, _ = solver.timing()
    solver.set_timing(False)
    peak, peak_src = measured_peak()
    per_gpu_cells = n * n
    ach_corr = per_gpu_cells * BYTES_CORR / (t_corr * 1e-3) / 1e9
    ach_pred = per_gpu_cells * BYTES_PRED / (t_pred * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": "fv_pass_kernel<corrector> (timed: the straight-line launch + the general launch that returns at once)", "achieved": ach_corr, "peak": peak,
                "unit": "GB/s", "frac": ach_corr / peak, "traffic": TRAFFIC_CORR_BYTES,
                "peak_source": peak_src, "algorithmic_bytes_per_launch": per_gpu_cells * BYTES_CORR,
                "ms_per_launch": t_corr,
                "predictor": {"achieved": ach_pred, "frac": ach_pred / peak, "ms_per_launch": t_pred,
                              "algorithmic_bytes_per_launch": per_gpu_cells * BYTES_PRED},
                "step_frac_of_hbm_roofline": (value / world) * BYTES_PER_CELL_UPDATE / 1e9 / peak}

    # ---- end to end through the host-buffer entry point ---------------------------------------
    # every step: H2D of this rank's q from pinned host memory, update(), D2H of q (+ scalars)
    qh = torch.empty((3, n + 2, n + 2), dtype=torch.float64, pin_memory=True)
    qh.copy_(torch.from_numpy(solver.get_q()))
    k2 = max(3, min(args.steps, 10))
    solver.update_host(qh, 1)
    barrier()
    t0 = time.perf_counter()
    for _ in range(k2):
        solver.update_host(qh, 1)            # synchronises
    barrier()
    t1 = time.perf_counter()
    t_e2e = t1 - t0
    if world > 1:
        t = torch.tensor([t_e2e], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_e2e = float(t.item())
    nbytes = qh.numel() * 8 * world
    e2e = {"value": cells_per_step * k2 / t_e2e, "unit": "cell-updates/s", "h2d_bytes_per_step": nbytes,
           "d2h_bytes_per_step": nbytes + (160 if world == 1 else 0), "steps": k2,
           "api": ("gpf_fv_update_host (C ABI; H2D / passes / D2H pipelined over 16 row chunks) via FVSolver.update_host" if world == 1 else
                   "SlabSolver.update_host (gpf_fv_upload / slab step / gpf_fv_download per rank)")
                  + ", pinned host q, all ranks' bytes"}

    # ---- the other arithmetic contract, same workload (N=1 only) -------------------------------
    other = None
    if world == 1:
        mode = "exact" if args.arith == "fast" else "fast"
        del solver
        torch.cuda.empty_cache()
        from gapflow_b200.fv import FVSolver
        s2 = FVSolver(d['grid'], d['geometry'], d['properties'], d['numerics'], topo=topo, arith=mode)
        s2.set_q(q0)
        s2.fill_ghosts()
        s2.pre_run(False)
        s2.step(args.warmup)
        torch.cuda.synchronize()
        k3 = max(3, args.steps // 2)
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        s2.step(k3)
        f1.record()
        torch.cuda.synchronize()
        ms2 = f0.elapsed_time(f1) / k3
        other = {"arith": mode, "value": n * n / (ms2 * 1e-3), "ms_per_step": ms2, "steps": k3,
                 "step_frac_of_hbm_roofline": n * n / (ms2 * 1e-3) * BYTES_PER_CELL_UPDATE / 1e9 / peak}
        del s2

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu:
            ncpu = 1024
            rate, _ = cpu_steps(ncpu, 3, 1)
            cpu = {"value": rate, "unit": "cell-updates/s", "cores": 1, "kind": "port",
                   "sample": f"{ncpu}x{ncpu} sample of the same slider, 3 steps after 1 warm-up, NumPy oracle "
                             f"(single-threaded like the reference); host has {os.cpu_count()} logical cores"}
        line = {
            "metric": "FV cell-updates/s", "value": value, "unit": "cell-updates/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"2D parabolic slider bearing {n * world}x{n} ({n}x{n} per GPU), FP64, DH EOS, "
                                   "fixed-form stress, x: D/N/N, y periodic, adaptive dt (BASELINE config 2)",
                       "parallelism": f"x-slabs x{world}" if world > 1 else "single GPU",
                       "arith": args.arith + (" (FMA contraction, a*rcp(b) quotients; parity 1e-12/step, 1e-9/1000 steps)"
                                              if args.arith == "fast" else " (bit-identical to the reference's NumPy path)"),
                       "l2": "inputs larger than L2: 1.2 GB of fields streamed per step vs 126 MB L2"},
            "clocks": clk.summary(), "host_enqueue_ms_per_step": 1e3 * t_host / args.steps,
            "gpu_launches": int(launches), "roofline": roofline,
            "e2e": e2e, "cpu_baseline": cpu, "other_arith": other,
            "final": {"step": int(sc.step), "dt": sc.dt, "residual": sc.residual},
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()


# dram__bytes_read.sum + dram__bytes_write.sum of one corrector launch at 4096^2 from the
# committed `ncu --set full` capture (profiles/); None until a capture exists.
TRAFFIC_CORR_BYTES = 1486998624  # profiles/r1s_ncu_fv_pass_summary.csv: 1.105212 GB read + 381.787 MB written


_RESULT_FD = None


def _claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version
    banner to fd 1 when NCCL_DEBUG is set on the box), so everything written to fd 1 during the run
    goes to stderr and the result line is written to the original stdout at the end."""
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    os.write(_RESULT_FD if _RESULT_FD is not None else 1, (json.dumps(line) + "\n").encode())


def main():
    _claim_stdout()
    # slab halo rows travel by copy engine, not by an NCCL kernel: they overlap the interior launch
    # of the pass, whose persistent CTAs leave no SM free (measured: -2 % ms/step at N = 2)
    os.environ.setdefault("NCCL_P2P_USE_CUDA_MEMCPY", "1")
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="gapflow_b200")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--arith", default="fast", choices=["fast", "exact"],
                    help="arithmetic contract of the pass kernels (include/gapflow_b200.h: gpf_fv_cfg.arith)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
