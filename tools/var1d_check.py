"""Device check of the opt-in O(log M) variance (GPF_GP_VAR1D=1) against the dense-solve kernel and the oracle.
usage: GPF_GP_VAR1D=1 python tools/var1d_check.py   (prints max |var - oracle| / amp, arg-max agreement)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from gapflow_b200.gp import GPModel
from oracle import gp as G
from tests.gp_cases import query_set, training_set

for M, model in ((300, 'press_wide'), (2000, 'press')):
    c = training_set(M, model)
    g = GPModel(M, 1, 1)
    g.factor(c['X'], c['Y'], c['log_amp'], c['log_scale'], c['Yerr'])
    xq = query_set(20000, c)
    m, v = g.repredict_mean_var(torch.as_tensor(xq, device='cuda'), 1.5e-8)
    L = G.build(c['X'], c['Yerr'], c['log_amp'], c['log_scale'])
    a = G.alpha(L, c['Y'])
    mr, vr = G.repredict_mean_var(c['X'], c['log_amp'], c['log_scale'], L, a, 1.5e-8, xq)
    v = v.cpu().numpy()
    print(f"VAR1D={os.environ.get('GPF_GP_VAR1D', '0')} M={M}: max|var-oracle|/amp {np.abs(v - vr).max():.3e}, "
          f"mean err {np.abs(m.cpu().numpy() - mr).max() / np.abs(mr).max():.3e}, "
          f"argmax {g.argmax} vs {int(np.argmax(vr))}, vmax {g._vmax_norm:.6e} vs {vr.max():.6e}")
    assert np.abs(v - vr).max() <= 1e-10 * np.exp(c['log_amp']) and g.argmax == int(np.argmax(vr))
print("VAR1D_CHECK_OK")
