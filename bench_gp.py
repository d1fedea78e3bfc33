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
    ap.add_argument("--var-sample", type=int, default=32768)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    import numpy as np
    import torch
    from gapflow_b200 import _capi
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
                        "kernel_evals_per_s": args.N * args.M / (t_mean * 1e-3)},
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
    print(json.dumps(out))


if __name__ == "__main__":
    main()
