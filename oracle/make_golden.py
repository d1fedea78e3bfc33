"""Generates tests/golden/*.npz from the REFERENCE's own functions -- TEST INFRASTRUCTURE ONLY.

Run in the build container (needs /root/reference):

    python -m oracle.make_golden

* fv_closures.npz : seeded random inputs and the outputs of the reference's
  eos_pressure / eos_sound_velocity (all 7 EOS) and stress_bottom / stress_top /
  stress_avg (slip="top", default dq) -- GaPFlow/models/{pressure,sound,viscous}.py.
* fv_steps.npz    : for several boundary-condition / MacCormack-order cases on small
  grids, the solution after 1 and 3 steps of the reference's update loop, where every
  arithmetic function (closures, predictor_corrector, source) is the reference's own,
  imported by file path; only the ~40-line driver and ghost-cell code of
  GaPFlow/problem.py (not importable as it is: needs muGrid/JAX) is the restatement in oracle/fv.py --
  which tests/test_oracle_vs_reference_problem.py holds to the reference's Problem class executed with stand-ins for
  those packages (oracle/ref_problem.py), together with this file's output.
"""
import os

import numpy as np

from . import fv, ref_loader

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")

EOS_PROPS = [
    dict(EOS='DH', rho0=877.7007, P0=101325., C1=3.5e10, C2=1.23),
    dict(EOS='PL', rho0=1.1853, P0=101325., alpha=0.3),
    dict(EOS='vdW', M=39.948, T=100., a=1.355, b=0.03201),
    dict(EOS='MT', rho0=700., P0=0.101e6, K=.557e9, n=7.33),
    dict(EOS='cubic', a=15.2, b=-9.6, c=3.35, d=-0.07),
    dict(EOS='BWR', T=2., gamma=3.),
    dict(EOS='Bayada', rho_l=850., rho_v=0.019, c_l=1600., c_v=352.),
]
EOS_DENS = {'DH': (877.7, 0.05), 'PL': (1.2, 0.1), 'vdW': (300., 0.1), 'MT': (700., 0.02),
            'cubic': (0.8, 0.1), 'BWR': (0.8, 0.1), 'Bayada': (450., 1.0)}


class RefDrivenOracle(fv.FVOracle):
    """FVOracle whose arithmetic is the reference's own functions."""
    R = None

    def closures(self):
        R, g, pr = self.R, self.geo, self.prop
        h = self.topo[:3]
        self.pressure[...] = R['pressure'].eos_pressure(self.q[0], pr)
        eta = self.viscosity()
        sb = R['viscous'].stress_bottom(self.q, h, g['U'], g['V'], eta, pr['bulk'], self.extra)
        st = R['viscous'].stress_top(self.q, h, g['U'], g['V'], eta, pr['bulk'], self.extra)
        for f, k in ((self.wall_xz, 4), (self.wall_yz, 3)):
            f[:3] = sb[:3] / 2.
            f[6:9] = st[:3] / 2.
            f[5] = sb[-1] / 2.
            f[11] = st[-1] / 2.
            f[k] = sb[k]
            f[k + 6] = st[k]
        self.bulk[...] = R['viscous'].stress_avg(self.q, h, g['U'], g['V'], eta, pr['bulk'], self.extra)

    def viscosity(self):
        """stress.py:305-326 with the reference's own viscosity.py functions."""
        V, g, pr = self.R['viscosity'], self.geo, self.prop
        mu0 = pr['shear']
        if 'piezo' in pr:
            mu0 = V.piezoviscosity(self.pressure if not pr['EOS'] == 'Bayada' else self.q[0], pr['shear'], pr['piezo'])
        if 'thinning' in pr:
            xx, yy = fv.midpoint_grid(self.grid)
            rate = V.shear_rate_avg(np.gradient(self.pressure, xx[:, 0], axis=0),
                                    np.gradient(self.pressure, yy[0, :], axis=1), self.topo[0], g['U'], g['V'], mu0)
            return mu0 * V.shear_thinning_factor(rate, mu0, pr['thinning'])
        return mu0

    def one_pass(self, d):
        R = self.R
        dx, dy, dt = self.grid['dx'], self.grid['dy'], self.dt
        self.closures()
        fX, fY = R['integrate'].predictor_corrector(self.q, self.pressure, self.bulk, d)
        src = R['integrate'].source(self.q, self.topo, self.bulk,
                                    self.wall_xz[:6] + self.wall_yz[:6], self.wall_xz[6:] + self.wall_yz[6:])
        self.q[...] = self.q - dt * (fX / dx + fY / dy - src)
        self.fill_ghosts()

    @property
    def v_sound(self):
        return self.R['sound'].eos_sound_velocity(self.q[0], self.prop).max()


def closure_vectors(seed=20251104, shape=(23, 17)):
    R = ref_loader.modules()
    rng = np.random.default_rng(seed)
    out = {}
    q = np.empty((3, *shape))
    q[0] = 877.7 * (1 + 1e-2 * rng.standard_normal(shape))
    q[1] = 40 * (1 + 0.3 * rng.standard_normal(shape))
    q[2] = 5 * rng.standard_normal(shape)
    h = np.empty((3, *shape))
    h[0] = 1e-5 * (1 + 0.5 * rng.random(shape))
    h[1] = 1e-3 * rng.standard_normal(shape)
    h[2] = 1e-3 * rng.standard_normal(shape)
    Ls = 1e-6 * rng.random((1, *shape))
    out['q'], out['h'], out['Ls'] = q, h, Ls[0]
    out['visc_params'] = np.array([[0.1, 0., 0.0794, 0.], [0.1, -0.03, 0.0794, 0.013]])   # U, V, eta, zeta
    for k, (U, V, eta, zeta) in enumerate(out['visc_params']):
        for ls_name, ls in (('ls0', np.zeros_like(Ls)), ('ls', Ls)):
            for fn in ('stress_bottom', 'stress_top', 'stress_avg'):
                out[f'{fn}_{k}_{ls_name}'] = getattr(R['viscous'], fn)(q, h, U, V, eta, zeta, ls)
    # the general signatures (viscous.py:37, 281, 612): slip at both walls, in-plane gradients
    rg = np.random.default_rng(seed + 1)           # its own generator: the draws below stay what they were
    out['dqx'] = np.stack([1e4 * rg.standard_normal(shape), 3e3 * rg.standard_normal(shape), 3e2 * rg.standard_normal(shape)])
    out['dqy'] = np.stack([1e4 * rg.standard_normal(shape), 3e2 * rg.standard_normal(shape), 3e3 * rg.standard_normal(shape)])
    for k, (U, V, eta, zeta) in enumerate(out['visc_params']):
        for slip in ('top', 'both'):
            for g_name, gx, gy in (('dq0', None, None), ('dq', out['dqx'], out['dqy'])):
                if slip == 'top' and g_name == 'dq0':
                    continue                       # the driver's call: covered bit for bit above
                for fn in ('stress_bottom', 'stress_top', 'stress_avg'):
                    out[f'{fn}_{k}_{slip}_{g_name}'] = getattr(R['viscous'], fn)(q, h, U, V, eta, zeta, Ls, dqx=gx, dqy=gy, slip=slip)
    for p in EOS_PROPS:
        mu, w = EOS_DENS[p['EOS']]
        d = mu * (1 + w * (rng.random(shape) - 0.5) * 2) if p['EOS'] != 'Bayada' else rng.random(shape) * 900.
        out[f"dens_{p['EOS']}"] = d
        out[f"p_{p['EOS']}"] = R['pressure'].eos_pressure(d.copy(), p)
        with np.errstate(invalid='ignore'):
            out[f"c_{p['EOS']}"] = R['sound'].eos_sound_velocity(d.copy(), p)
    return out


def step_cases():
    """name -> (yaml text, seed, extra?)"""
    from tests.cases import STEP_CASES
    return STEP_CASES


def step_vectors():
    from tests.cases import build_case
    RefDrivenOracle.R = ref_loader.modules()
    out = {}
    for name in step_cases():
        o = build_case(name, cls=RefDrivenOracle)
        out[f'{name}_q0'] = o.q.copy()
        out[f'{name}_dt0'] = np.array(o.dt)
        o.update()
        out[f'{name}_q1'] = o.q.copy()
        out[f'{name}_s1'] = np.array([o.dt, o.residual, o.kinetic_energy, o.v_max, o.v_sound, o.mass])
        o.update()
        o.update()
        out[f'{name}_q3'] = o.q.copy()
        out[f'{name}_s3'] = np.array([o.dt, o.residual, o.kinetic_energy, o.v_max, o.v_sound, o.mass])
    return out


def main():
    assert ref_loader.available(), "reference not found"
    os.makedirs(GOLDEN, exist_ok=True)
    np.savez_compressed(os.path.join(GOLDEN, "fv_closures.npz"), **closure_vectors())
    np.savez_compressed(os.path.join(GOLDEN, "fv_steps.npz"), **step_vectors())
    print("wrote", os.listdir(GOLDEN))


if __name__ == "__main__":
    main()
