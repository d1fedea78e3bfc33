#!/usr/bin/env python
"""GP inference benchmark (BASELINE.json metric, second half: "GP predictions/s").

Workload = BASELINE config 4: the 2-D GP surrogate at 1024 x 1024 cells (N = 1026^2 = 1 052 676 query
cells incl. ghosts) against M = 10 000 training points; three models as in
examples/slip_2d_lj_mock.py: pressure (D = 1, n_out = 1), shear-x and shear-y (D = 3, n_out = 2).
One prediction = one query cell evaluated by one model.  Reports, per model: factorisation time,
mean predictions/s (full N), mean+variance predictions/s (on a query subsample, stated), FP64 rates
against a cuBLAS DGEMM peak measured in the same run, and the SciPy oracle on the host cores.

    python bench_gp.py [--M 10000] [--N 1052676] [--var-sample 32768] [--no-cpu]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def dgemm_peak(torch):
    n = 8192
    a = torch.randn(n, n, dtype=torch.float64, device='cuda')
    b = torch.randn(n, n, dtype=torch.float64, device='cuda')
    torch.matmul(a, b)
    best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return 2.0 * n**3 / (best * 1e-3) / 1e12


def timed(torch, fn, reps=1):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--M", type=int, default=10000)
    ap.add_argument("--N", type=int, default=1026 * 1026)
    ap.add_argument("--var-sample", type=int, default=75776)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--coupled-n", type=int, default=1024, help="cells per side of the GP-coupled FV step (0: skip)")
    ap.add_argument("--coupled-steps", type=int, default=3)
    ap.add_argument("--coupled-only", action="store_true", help="only the GP-coupled step (profiling runs)")
    args = ap.parse_args()
    import numpy as np
    import torch
    from gapflow_b200 import _capi
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        # torchrun: the GP-coupled step on `world` slabs (training set replicated, query cells partitioned)
        import torch.distributed as dist
        torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
        dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
        args.no_cpu = True
        res = coupled_step(torch, np, _capi.load(), args, world=world)
        if dist.get_rank() == 0:
            print(json.dumps({"metric": "GP-coupled FV cell-updates/s", "unit": "cell-updates/s", "n_gpus": world,
                              "value": res["cell_updates_per_s"], "scaling": "strong", "dtype": "f64", "data": "synthetic",
                              "coupled_step": res}))
        dist.barrier()
        dist.destroy_process_group()
        return
    if args.coupled_only:
        print(json.dumps({"metric": "GP-coupled FV cell-updates/s", "coupled_step": coupled_step(torch, np, _capi.load(), args)}))
        return
    from gapflow_b200.gp import GPModel
    from tests.gp_cases import query_set, training_set
    assert torch.cuda.is_available(), "needs a CUDA device: there is no CPU path"
    lib = _capi.load()
    peak = dgemm_peak(torch)
    out = {"metric": "GP predictions/s", "unit": "predictions/s", "n_gpus": 1, "dtype": "f64", "data": "synthetic",
           "config": {"workload": f"2D GP surrogate, N = {args.N} query cells x M = {args.M} training points "
                                  "(BASELINE config 4), Matern-3/2, mock LJ training data (seed 2025)"},
           "fp64_peak_tflops": peak, "fp64_peak_source": "cuBLAS DGEMM 8192^3 via torch.matmul, best of 5, same run",
           "models": {}}
    for name, model in (("pressure", "press"), ("shear_x", "shear_x")):
        c = training_set(args.M, model)
        D, nout = len(c['dims']), c['n_out']
        g = GPModel(args.M, D, nout)
        l0 = lib.gpf_launch_count()
        t_fac = timed(torch, lambda: g.factor(c['X'], c['Y'], c['log_amp'], c['log_scale'], c['Yerr']))
        xt = torch.as_tensor(query_set(args.N, c), device='cuda')
        t_mean = timed(torch, lambda: g.repredict_mean(xt), reps=3)
        ns = min(args.var_sample, args.N)
        xs = xt[:ns].contiguous()
        t_mv = timed(torch, lambda: g.repredict_mean_var(xs, 1.5e-8))
        t_m_s = timed(torch, lambda: g.repredict_mean(xs), reps=3)
        # one evaluation of the training objective and its gradient (gp.py:306-318) after a factorisation
        t_nlml = timed(torch, lambda: g.nlml())
        launches = lib.gpf_launch_count() - l0
        Mp = (args.M + 63) // 64 * 64
        flop_var = float(Mp) * Mp * ns            # forward substitution: M^2/2 multiply-adds per query
        flop_fac = args.M**3 / 3.0
        res = {"D": D, "n_out": nout,
               "factor_ms": t_fac, "factor_tflops": flop_fac / (t_fac * 1e-3) / 1e12,
               "nlml_grad_ms": t_nlml, "nlml_grad_tflops": 2.0 * flop_fac / (t_nlml * 1e-3) / 1e12,
               "mean": {"value": args.N / (t_mean * 1e-3), "ms": t_mean, "N": args.N,
                        "algorithm": ("O(log M) per query: sorted training points, blockwise running sums (csrc/gp_mean1d.h)"
                                      if D == 1 and nout == 1 and os.environ.get("GPF_GP_MEAN1D", "1") != "0"
                                      else "O(M) per query: direct sum over the training points"),
                        "equivalent_kernel_evals_per_s": args.N * args.M / (t_mean * 1e-3)},
               "mean_var": {"value": ns / (t_mv * 1e-3), "ms": t_mv, "N": ns,
                            "sample": f"first {ns} of the {args.N} query cells (the variance costs M^2 flops per cell)",
                            "var_kernel_tflops": flop_var / ((t_mv - t_m_s) * 1e-3) / 1e12,
                            "frac_of_dgemm_peak": flop_var / ((t_mv - t_m_s) * 1e-3) / 1e12 / peak},
               "gpu_launches": int(launches)}
        if not args.no_cpu and name == "pressure":
            from oracle import gp as G
            nc = 4096
            xh = xs[:nc].cpu().numpy()
            t0 = time.perf_counter()
            L = G.build(c['X'], c['Yerr'], c['log_amp'], c['log_scale'])
            a = G.alpha(L, c['Y'])
            t1 = time.perf_counter()
            G.repredict_mean(c['X'], c['log_amp'], c['log_scale'], a, xh)
            t2 = time.perf_counter()
            G.repredict_mean_var(c['X'], c['log_amp'], c['log_scale'], L, a, 1.5e-8, xh)
            t3 = time.perf_counter()
            res["cpu_baseline"] = {"kind": "port", "cores": os.cpu_count(), "unit": "predictions/s",
                                   "sample": f"{nc} query cells, SciPy restatement of gp.py:509-535 (OpenBLAS, all host cores)",
                                   "factor_s": t1 - t0, "mean": nc / (t2 - t1), "mean_var": nc / (t3 - t2)}
        out["models"][name] = res
        del g
    out["value"] = out["models"]["pressure"]["mean"]["value"]
    if args.coupled_n > 0:
        out["coupled_step"] = coupled_step(torch, np, lib, args)
    print(json.dumps(out))


COUPLED_YAML = """
options: {{output: /tmp/gpf_bench, write_freq: 1000000, use_tstamp: False, silent: True}}
grid: {{Lx: 1470., Ly: 1470., Nx: {n}, Ny: {n}, xE: ['P','P','P'], xW: ['P','P','P'], yS: ['P','P','P'], yN: ['P','P','P']}}
geometry: {{type: journal, hmin: 10., hmax: 14., U: 0.12, V: 0.02}}
numerics: {{CFL: 0.4, adaptive: 1, tol: 1e-8, dt: 0.1, max_it: 1000000}}
properties: {{shear: 2.15, bulk: 0.5, EOS: BWR, T: 1.0, rho0: 0.8}}
gp:
  press: {{atol: 1., rtol: 0., obs_stddev: 2.e-2, active_learning: False, active_dims: [0]}}
  shear: {{atol: 1., rtol: 0., obs_stddev: 4.e-3, active_learning: False, active_dims: {{x: [0, 1, 6], y: [0, 2, 6]}}}}
db: {{init_size: {M}, init_method: lhc, init_width: 0.02}}
"""


def _coupled_inputs(np, n, M):
    from gapflow_b200 import config
    d = config.read_yaml_string(COUPLED_YAML.format(n=n, M=M))
    i = (np.arange(n + 2) - 0.5) / n
    extra = (0.5 + 0.4 * np.sin(2 * np.pi * i)[:, None] * np.cos(2 * np.pi * i)[None, :])[None]
    return d, extra


def _fixed_hyper(model):
    import numpy as np
    return 0.0, np.log(np.std(np.asarray(model.Xtrain), axis=0))       # params_init (stress.py:281-284)


def coupled_step(torch, np, lib, args, world=1):
    """One Problem.update() with all three GP models (slip_2d_lj_mock.py scaled: BASELINE config 4): per pass the
    feature matrices, three predictive means over every padded cell and the surrogate pass kernel; then the
    GP sound speed.  Active learning off, hyperparameters = params_init (no optimiser in the timed region)."""
    from gapflow_b200.fv import FVSolver
    from gapflow_b200.surrogate import Database, Mock, SurrogateDriver
    n, M = args.coupled_n, args.M
    d, extra = _coupled_inputs(np, n, M)
    from gapflow_b200.topography import make_topography
    topo = make_topography(d['grid'], d['geometry'])
    db = Database(Mock(d['properties'], d['geometry'], d['gp'], draws=(0.3, -0.3, 0.5)), d['db'])
    if world > 1:
        import torch.distributed as dist
        from gapflow_b200.dist import SlabSolver
        from gapflow_b200.surrogate import DistColl, SlabSurrogateDriver
        sl = SlabSolver.from_global(d, None, topo, extra, arith='fast')
        s = sl.fv
        drv = SlabSurrogateDriver(sl, d['gp'], db, 2, DistColl(), write_freq=10**9, optimizer=_fixed_hyper)
    else:
        s = FVSolver(d['grid'], d['geometry'], d['properties'], d['numerics'], topo=topo, extra=extra, arith='fast')
        drv = SurrogateDriver(s, d['gp'], db, 2, write_freq=10**9, optimizer=_fixed_hyper)
    t0 = time.perf_counter()
    drv.pre_run()                                # initial design (M mock-MD rows), 3 factorisations, first inference
    torch.cuda.synchronize()
    t_pre = time.perf_counter() - t0
    l0 = lib.gpf_launch_count()
    drv.update(0)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(args.coupled_steps):
        drv.update(1 + k)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.coupled_steps
    if world > 1:                                  # device time, max over ranks
        t = torch.tensor([ms], dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    launches = (lib.gpf_launch_count() - l0) // (args.coupled_steps + 1)
    sc = s.scalars()
    Np = (n + 2) * (n + 2)
    res = {"workload": f"2D journal film {n}x{n}, BWR fluid, pressure + shear-x + shear-y GP (D = 1, 3, 3) on M = {db.size} "
                       "mock-MD points, active learning off",
           "ms_per_step": ms, "cell_updates_per_s": n * n / (ms * 1e-3),
           "gp_predictions_per_step": 7 * Np, "gp_predictions_per_s": 7 * Np / (ms * 1e-3),
           "equivalent_kernel_evals_per_s": 7.0 * Np * db.size / (ms * 1e-3),
           "pre_run_s": t_pre, "gpu_launches_per_step": int(launches), "final": {"step": int(sc.step), "dt": sc.dt},
           "n_gpus": world, "parallelism": "single GPU" if world == 1 else
           f"{world} x-slabs: query cells partitioned, training set / factor replicated, halo rows over peer memory"}
    if not args.no_cpu:
        from oracle import surrogate as OS
        nc = 104                                      # (nc + 2)^2 >= M: the design picks M distinct cells (db.py:329)
        dc, ec = _coupled_inputs(np, nc, M)
        odb = OS.Database(OS.MockMD(dc['properties'], dc['geometry'], dc['gp'], draws=(0.3, -0.3, 0.5)), dc['db'])
        o = OS.SurrogateFVOracle(dc['grid'], dc['geometry'], dc['properties'], dc['numerics'], dc['gp'], odb, extra=ec,
                                 write_freq=10**9, optimizer=_fixed_hyper)
        o.pre_run()
        t0 = time.perf_counter()
        o.update()
        t1 = time.perf_counter()
        res["cpu_baseline"] = {"kind": "port", "cores": os.cpu_count(), "unit": "cell-updates/s",
                               "value": nc * nc / (t1 - t0),
                               "sample": f"one step of the NumPy/SciPy oracle (oracle/surrogate.py) on {nc}x{nc} cells, same M"}
    return res


if __name__ == "__main__":
    main()
