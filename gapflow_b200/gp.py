"""Device GP surrogate: host mirror of the reference's inference interface over the C ABI.

``GPModel`` stands in for the tinygp object + cache that ``GaussianProcessSurrogate`` keeps
(GaPFlow/models/gp.py:268-288, 337-414, 509-570): ``factor`` = build_gp + the solves of
``gp.condition``; ``repredict_mean`` / ``repredict_mean_var`` = the jitted free functions of the
same names.  Inputs and outputs are in the reference's NORMALISED units; ``infer`` applies
``Yscale`` like ``_infer_mean`` / ``_infer_mean_var`` (gp.py:362, 384-386) and keeps
``maximum_variance`` / ``variance_tol`` (gp.py:406-410).  Hyperparameters are inputs: the
marginal-likelihood optimiser stays on the host.
"""
import ctypes as C

import numpy as np
import torch

from . import _capi


class GPModel:
    def __init__(self, M_max, D, n_out=1, device=None, arith=None):
        if not torch.cuda.is_available():
            raise _capi.GapflowLibraryError("gapflow_b200 needs a CUDA device; there is no CPU path")
        self.lib = _capi.load()
        self.device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
        torch.cuda.set_device(self.device)
        self.M_max, self.D, self.n_out = int(M_max), int(D), int(n_out)
        nbytes = C.c_size_t()
        _capi.check(self.lib.gpf_gp_workspace_bytes(self.M_max, self.D, self.n_out, C.byref(nbytes)))
        self.workspace = torch.empty(nbytes.value, dtype=torch.uint8, device=self.device)
        h = C.c_void_p()
        _capi.check(self.lib.gpf_gp_create(self.M_max, self.D, self.n_out, C.c_void_p(self.workspace.data_ptr()),
                                           nbytes.value, C.byref(h)))
        self.h = h
        if arith is not None:                 # 'fast' (library default) | 'exact': include/gapflow_b200.h gpf_gp_set_arith
            _capi.check(self.lib.gpf_gp_set_arith(self.h, _capi.ARITH[arith]))
        self.M = 0
        self.log_amp, self.log_scale, self.yerr = None, None, None
        self.maximum_variance, self.variance_tol, self.argmax = None, None, None

    def __del__(self):
        h = getattr(self, 'h', None)
        if h:
            self.lib.gpf_gp_destroy(h)
            self.h = None

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # -- build_gp + condition (gp.py:268-288, 546-568) ------------------------------------
    def factor(self, Xtrain, Y, log_amp, log_scale, yerr):
        X = np.ascontiguousarray(Xtrain, dtype=np.float64).reshape(-1, self.D)
        Y = np.ascontiguousarray(np.asarray(Y, dtype=np.float64).reshape(len(X), self.n_out))
        ls = np.ascontiguousarray(np.broadcast_to(np.asarray(log_scale, dtype=np.float64), (self.D,)))
        _capi.check(self.lib.gpf_gp_factor(self.h, C.c_void_p(X.ctypes.data), len(X), float(log_amp),
                                           C.c_void_p(ls.ctypes.data), float(yerr), C.c_void_p(Y.ctypes.data),
                                           self._stream()))
        self.M, self.log_amp, self.log_scale, self.yerr = len(X), float(log_amp), ls.copy(), float(yerr)

    # -- training (gp.py:290-335) -----------------------------------------------------------
    def nlml(self):
        """(loss, gradient w.r.t. (log_amp, log_scale[0..D-1])) at the parameters of the last
        ``factor``: the objective of ``_train`` and what JAX autodiff hands its optimiser."""
        nb = C.c_size_t()
        _capi.check(self.lib.gpf_gp_train_workspace_bytes(self.M, C.byref(nb)))
        if getattr(self, '_train_ws', None) is None or self._train_ws.numel() < nb.value:
            self._train_ws = torch.empty(nb.value, dtype=torch.uint8, device=self.device)
        loss = C.c_double()
        grad = (C.c_double * (1 + self.D))()
        _capi.check(self.lib.gpf_gp_nlml(self.h, C.c_void_p(self._train_ws.data_ptr()), self._train_ws.numel(),
                                         C.byref(loss), grad, self._stream()))
        return loss.value, np.array(grad[:])

    def train(self, Xtrain, Y, yerr, log_amp0=0.0, log_scale0=None):
        """``_train``: minimise the negative log marginal likelihood from ``params_init``
        (stress.py:281-284: log_amp = 0, log_scale = log(std(Xtrain, axis=0))) with SciPy's BFGS -- the
        default method behind jaxopt.ScipyMinimize -- every evaluation a device factorisation plus
        the device gradient; ends with the final build (factor) at the optimum.
        -> (log_amp, log_scale, loss)"""
        from scipy.optimize import minimize
        X = np.ascontiguousarray(Xtrain, dtype=np.float64).reshape(-1, self.D)
        if log_scale0 is None:
            log_scale0 = np.log(np.std(X, axis=0))
        x0 = np.concatenate([[float(log_amp0)], np.asarray(log_scale0, dtype=np.float64).reshape(self.D)])

        def fun(t):
            try:
                self.factor(X, Y, t[0], t[1:], yerr)
            except _capi.GapflowError:          # not positive definite at a trial point of the line search
                return np.inf, np.zeros_like(t)
            return self.nlml()

        res = minimize(fun, x0, jac=True, method='BFGS')
        self.factor(X, Y, res.x[0], res.x[1:], yerr)
        return float(res.x[0]), res.x[1:].copy(), float(res.fun)

    def get_factor(self):
        """(L (M, M) lower, alpha (M, n_out)) as host arrays."""
        L = np.empty((self.M, self.M))
        a = np.empty((self.M, self.n_out))
        _capi.check(self.lib.gpf_gp_get_factor(self.h, C.c_void_p(L.ctypes.data), C.c_void_p(a.ctypes.data), self._stream()))
        return L, a

    # -- re-predict functions (gp.py:509-535) ---------------------------------------------
    def _x(self, Xtest):
        x = Xtest if isinstance(Xtest, torch.Tensor) else torch.as_tensor(np.ascontiguousarray(Xtest), device=self.device)
        x = x.to(torch.float64).reshape(-1, self.D).contiguous()
        assert x.is_cuda
        return x

    def repredict_mean(self, Xtest, with_gradient_max=False):
        """-> mean (n_out, N) [, max_i d mean_0 / d x_0]"""
        x = self._x(Xtest)
        N = x.shape[0]
        mean = torch.empty((self.n_out, N), dtype=torch.float64, device=self.device)
        g = C.c_double()
        _capi.check(self.lib.gpf_gp_mean(self.h, C.c_void_p(x.data_ptr()), N, C.c_void_p(mean.data_ptr()),
                                         C.byref(g) if with_gradient_max else None, self._stream()))
        return (mean, g.value) if with_gradient_max else mean

    def repredict_mean_var(self, Xtest, noise=0.0):
        """-> mean (n_out, N), var (N); sets maximum_variance (normalised) and argmax."""
        x = self._x(Xtest)
        N = x.shape[0]
        mean = torch.empty((self.n_out, N), dtype=torch.float64, device=self.device)
        var = torch.empty(N, dtype=torch.float64, device=self.device)
        nptr, nsc = None, 0.0
        if isinstance(noise, torch.Tensor):
            noise = noise.to(torch.float64).contiguous()
            nptr = C.c_void_p(noise.data_ptr())
        else:
            nsc = float(noise)
        vmax, imax = C.c_double(), C.c_int64()
        _capi.check(self.lib.gpf_gp_mean_var(self.h, C.c_void_p(x.data_ptr()), N, nptr, nsc, C.c_void_p(mean.data_ptr()),
                                             C.c_void_p(var.data_ptr()), C.byref(vmax), C.byref(imax), self._stream()))
        self._vmax_norm, self.argmax = vmax.value, imax.value
        return mean, var

    # -- _infer (gp.py:388-414) -------------------------------------------------------------
    def infer(self, Xtest, shape, Yscale, Yerr, atol, rtol, compute_var=True, noise=0.0):
        """predictive mean (and variance) reshaped to the field shape and scaled like the reference."""
        if compute_var:
            m, v = self.repredict_mean_var(Xtest, noise)
            self.maximum_variance = self._vmax_norm * Yscale**2
            self.variance_tol = max(atol * Yerr * Yscale, rtol * Yscale)**2
            return (m.reshape(-1, *shape).squeeze(0) if self.n_out == 1 else m.reshape(-1, *shape)) * Yscale, \
                v.reshape(*shape) * Yscale**2
        m = self.repredict_mean(Xtest)
        return (m.reshape(-1, *shape).squeeze(0) if self.n_out == 1 else m.reshape(-1, *shape)) * Yscale, None


def features(fv_solver, X_scale, active_dims):
    """(Np, D) normalised test inputs of the FV solver's current solution (gp.py:223-232,
    stress.py:194-197), built on the device."""
    lib = _capi.load()
    dims = (C.c_int32 * len(active_dims))(*[int(d) for d in active_dims])
    xs = np.ascontiguousarray(np.asarray(X_scale, dtype=np.float64).reshape(7))
    Np = fv_solver.shape[0] * fv_solver.shape[1]
    X = torch.empty((Np, len(active_dims)), dtype=torch.float64, device=fv_solver.device)
    _capi.check(lib.gpf_gp_features(fv_solver.h, C.c_void_p(xs.ctypes.data), dims, len(active_dims),
                                    C.c_void_p(X.data_ptr()), fv_solver._stream()))
    return X
