"""Short GP workload for ncu (BASELINE config 4 shapes): one factorisation at M = 10^4, the dense variance on
one wave of query tiles (296 x 128 cells), the D = 3 two-output mean on 2^18 cells, one training objective.
Order of the kernels: build_K, (potrf_diag, panel, syrk) x 157, trsv x 314, mean, var, argmax, mean (2^18),
trtri, lauum, grad, nlml_finish.   python tools/gp_profile.py [M]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from gapflow_b200.gp import GPModel

M = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
rng = np.random.default_rng(2025)
X = np.column_stack([rng.uniform(0.98, 1.0, M), rng.uniform(0.33, 1.0, M), rng.uniform(0., 1., M)])
Y = np.column_stack([np.sin(6 * X[:, 0]) + X[:, 1] * X[:, 2], np.cos(5 * X[:, 1]) - X[:, 2]]) + 4e-3 * rng.standard_normal((M, 2))
ls = np.log(np.std(X, axis=0))
g = GPModel(M, 3, 2)
g.factor(X, Y, 0.0, ls, 4e-3)
xt = torch.as_tensor(rng.uniform(X.min(0), X.max(0), (1 << 18, 3)), device='cuda')
m, v = g.repredict_mean_var(xt[:296 * 128].contiguous(), 1.5e-8)      # one full wave: 2 CTAs per SM x 148 SMs x 128 query cells
mm = g.repredict_mean(xt)
loss, grad = g.nlml()
torch.cuda.synchronize()
print("var range", float(v.min()), float(v.max()), "loss", loss, "mean[0]", float(mm[0, 0]))
