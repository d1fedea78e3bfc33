"""Times gpf_gp_factor at M training points (for an ncu launch list). usage: gp_factor_probe.py [M] [reps]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gapflow_b200.gp import GPModel
M = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
rng = np.random.default_rng(1)
X = rng.uniform(0.1, 1.0, (M, 3)); Y = np.sin(3 * X[:, :1]) + 0.1 * X[:, 1:2]
gp = GPModel(M, 3, 1)
for r in range(reps):
    torch.cuda.synchronize(); t = time.perf_counter()
    gp.factor(X, Y, log_amp=0.1, log_scale=np.array([-0.5, 0.2, 0.1]), yerr=1e-3)
    torch.cuda.synchronize(); print("factor M=%d: %.2f ms" % (M, 1e3 * (time.perf_counter() - t)))
