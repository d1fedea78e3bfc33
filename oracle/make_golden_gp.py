"""Generates tests/golden/gp_coupled.npz from the EXECUTED reference -- TEST INFRASTRUCTURE ONLY.

Run in the build container (needs /root/reference):

    python -m oracle.make_golden_gp

The reference's own Problem / GaussianProcessSurrogate / Pressure / WallStress / Database / Mock classes run the
GP-coupled step (oracle/ref_problem.py) with oracle/tinygp_stand_in.py in tinygp's place -- NOT tinygp: GP parity stays
unpinned at that boundary -- and with the explicit inputs named in tests/surrogate_cases.py (hyper-parameters as a function
of the training inputs, the mock MD's three draws, the seeded design).  Per case and step the file holds the solution,
dt, the pressure and wall-stress fields, the variance fields, the database size; at the end the database and the
hyper-parameter histories.  tests/test_surrogate_golden.py holds oracle/surrogate.py to it wherever the tests run
(the reference tree is absent on the GPU box); the GPU parity tests compare the CUDA path with that oracle.
"""
import contextlib
import io
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")
FIELD = {'zz': 'pressure', 'xz': 'wall_stress_xz', 'yz': 'wall_stress_yz'}


def reference_run(ref, name, steps, al=True, **kw):
    """-> dict of arrays: the executed reference's GP-coupled run of a case of tests/surrogate_cases.py"""
    from tests import surrogate_cases as SC
    from . import tinygp_stand_in as T
    d, text = SC.inputs(name, **kw)
    g = d['grid']
    extra = SC.slip_length(g['Nx'], g['Ny'])
    with io.StringIO(text) as f, contextlib.redirect_stdout(io.StringIO()):
        rd = ref.read_yaml_input(f)
    rd['options']['silent'] = True
    rd['db']['dtool_path'] = None
    if not al:
        for k in ('press', 'shear'):
            rd['gp'][k]['active_learning'] = False
    T.HOOKS['draws'] = SC.DRAWS
    T.HOOKS['optimizer'] = lambda p0, X=None, **k: dict(zip(('log_amp', 'log_scale'),
                                                            SC.fixed_optimizer(type('M', (), {'Xtrain': X}))))
    out = {}
    with contextlib.redirect_stdout(io.StringIO()):
        db = ref.Database(ref.Mock(rd['properties'], rd['geometry'], rd['gp']), rd['db'])
        p = ref.Problem(rd['options'], rd['grid'], rd['numerics'], rd['properties'], rd['geometry'], gp=rd['gp'],
                        database=db, extra_field=extra)
        p.q[...] = SC.perturb(p.q)
        p._communicate_ghost_buffers()
        p._pre_run()
        models = {k: getattr(p, f) for k, f in FIELD.items() if getattr(p, f).is_gp_model}

        def snap(tag):
            out[f'q_{tag}'] = p.q.copy()
            out[f's_{tag}'] = np.array([p.dt, p.residual, float(p.pressure.v_sound), db.size, p.step])
            out[f'p_{tag}'] = np.array(p.pressure.pressure)
            out[f'wxz_{tag}'] = np.array(p.wall_stress_xz.full)
            out[f'wyz_{tag}'] = np.array(p.wall_stress_yz.full)
            for k, m in models.items():
                out[f'var_{k}_{tag}'] = np.array(m.variance)
                out[f'm_{k}_{tag}'] = np.array([float(m.maximum_variance), float(m.variance_tol), m._pause,
                                                m.last_fit_train_size, float(m.kernel_variance), float(m.Yscale), float(m.Yerr)])
        snap('pre')
        for k in range(steps):
            p.update()
            snap(str(k))
    out['X'], out['Y'], out['Yerr'] = np.array(db._Xtrain), np.array(db._Ytrain), np.array(db._Ytrain_err)
    for k, m in models.items():
        for key, h in m.history.items():
            out[f'hist_{k}_{key}'] = np.array([np.asarray(x, dtype=float) for x in h], dtype=float)
    return out


CASES = {       # name in the file -> (case of tests/surrogate_cases.py, keyword arguments, steps, active learning)
    'slip1d': ('slip1d', dict(nx=100, atol=0.6, al_steps=3), 8, True),
    'slip2d': ('slip2d', dict(nx=24, ny=20, atol=0.8, al_steps=2), 4, True),
    'slip1d_pause': ('slip1d', dict(nx=60, atol=0.05, al_steps=2), 6, True),
    'slip1d_noal': ('slip1d', dict(nx=60, wf=3), 7, False),
}


def main():
    from . import ref_problem
    ref = ref_problem.reference_package()
    out = {}
    for tag, (name, kw, steps, al) in CASES.items():
        for k, v in reference_run(ref, name, steps, al=al, **kw).items():
            out[f'{tag}/{k}'] = v
    path = os.path.join(GOLDEN, "gp_coupled.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, len(out), "arrays,", os.path.getsize(path), "bytes")
    ref_problem.release()


if __name__ == "__main__":
    main()
