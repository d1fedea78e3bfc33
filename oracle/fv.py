"""NumPy restatement of GaPFlow's finite-volume step -- TEST INFRASTRUCTURE ONLY.

Follows (reference file:line):

* GaPFlow/problem.py:509-569   Problem.update          -> FVOracle.update
* GaPFlow/problem.py:571-586   Problem._post_update    -> FVOracle._post_update
* GaPFlow/problem.py:588-610   Problem._finalize       -> FVOracle._finalize
* GaPFlow/problem.py:676-768   ghost cells             -> FVOracle.fill_ghosts / ghost_values
* GaPFlow/problem.py:314-362   scalar properties       -> FVOracle.{mass,kinetic_energy,...}
* GaPFlow/problem.py:662-670   _initialize             -> FVOracle.__init__
* GaPFlow/integrate.py:38-198  fluxes and source       -> predictor_corrector / source
* GaPFlow/models/stress.py:289-362, 427-459, 600-622   closure scatter -> FVOracle.closures
* GaPFlow/topography.py:38-255 grid + gap geometries   -> midpoint_grid / make_topography

Arrays have the reference's shapes: (ncomp, Nx+2, Ny+2), C order, float64.

PINNED: bit for bit to the golden vectors written from the reference's own functions (oracle/make_golden.py,
tests/test_oracle_golden.py), and -- where the reference tree is present -- to the reference's own ``Problem`` class
EXECUTED next to it (oracle/ref_problem.py; tests/test_oracle_vs_reference_problem.py: every golden case, ghost cells
on garbage, the invalid-state exit, the run loop; tests/test_oracle_vs_reference_random.py: seeded random configurations),
and it passes the reference's known-answer tests (tests/test_oracle_known_answers.py).
"""
from collections import deque
import numpy as np

from . import closures as C


# ----------------------------------------------------------------------------
# topography.py
# ----------------------------------------------------------------------------

def midpoint_grid(grid):
    """topography.py:38-54."""
    Nx, Ny, Lx, Ly = grid['Nx'], grid['Ny'], grid['Lx'], grid['Ly']
    x = np.arange(-1, Nx + 1) / Nx * Lx + (Lx / Nx) / 2.
    y = np.arange(-1, Ny + 1) / Ny * Ly + (Ly / Ny) / 2.
    return np.meshgrid(x, y, indexing='ij')


def make_topography(grid, geo):
    """topography.py:57-170 (generators) and :227-255 (flip, field fill).
    Returns (4, Nx+2, Ny+2): h, dh/dx, dh/dy, deformation(=0)."""
    xx, yy = midpoint_grid(grid)
    Lx = grid['Lx']
    t = geo['type']
    if t == 'journal':
        freq = 2. * np.pi / Lx
        if 'eps' in geo.keys():           # reference: `'CR' and 'eps' in keys`
            shift = geo['CR'] / freq
            amp = geo['eps'] * shift
        else:
            amp = (geo['hmax'] - geo['hmin']) / 2.
            shift = (geo['hmax'] + geo['hmin']) / 2.
        h = shift + amp * np.cos(freq * xx)
        hx = -amp * freq * np.sin(freq * xx)
        hy = np.zeros_like(h)
    elif t == 'inclined':
        slope = (geo['hmin'] - geo['hmax']) / Lx
        h = geo['hmax'] + slope * xx
        hx = np.ones_like(h) * slope
        hy = np.zeros_like(h)
    elif t == 'parabolic':
        pre = 4. / Lx**2 * (geo['hmax'] - geo['hmin'])
        h = pre * (xx - Lx / 2.)**2 + geo['hmin']
        hx = 2 * pre * (xx - Lx / 2.)
        hy = np.zeros_like(h)
    elif t == 'cdc':
        h0, h1, b = geo['hmin'], geo['hmax'], geo['b']
        slope = (h1 - h0) / (Lx / 2 - 2 * b)
        conv = np.logical_and(xx >= b, xx < Lx / 2 - b)
        cen = np.logical_and(xx >= Lx / 2 - b, xx < Lx / 2 + b)
        div = np.logical_and(xx >= Lx / 2 + b, xx < Lx - b)
        h = np.ones_like(xx) * h1
        h[conv] = h1 - slope * (xx[conv] - b)
        h[cen] = h0
        h[div] = h0 + slope * (xx[div] - (Lx / 2 + b))
        hx = np.zeros_like(h)
        hy = np.zeros_like(h)
        hx[conv] = -slope
        hx[div] = slope
    elif t == 'asperity':
        h, hx, hy = _asperity(xx, yy, grid, geo)
    else:
        raise ValueError(t)
    out = np.zeros((4, *h.shape))
    if geo.get('flip', False):
        out[0], out[2], out[1] = h.T, hx.T, hy.T
    else:
        out[0], out[1], out[2] = h, hx, hy
    return out


def _asperity(xx, yy, grid, geo):
    """topography.py:125-170, num == 1 only (num > 1 draws unseeded random heights)."""
    h0, h1, num = geo['hmin'], geo['hmax'], geo.get('num', 1)
    if num != 1:
        raise ValueError("asperity with num > 1 is not reproducible (unseeded RNG in the reference)")
    Lx, Ly = grid['Lx'], grid['Ly']
    m = np.logical_and((xx // (Lx / num)).astype(int) == 0, (yy // (Ly / num)).astype(int) == 0)
    bx, by = np.pi / (Lx / num), np.pi / (Ly / num)
    h = np.ones_like(xx) * h1
    hx = np.zeros_like(h)
    hy = np.zeros_like(h)
    cx, cy = np.mean(xx[m]), np.mean(yy[m])
    h[m] -= (h1 - h0) * (np.cos(bx * (xx[m] - cx)) * np.cos(by * (yy[m] - cy)))
    hx[m] += bx * (h1 - h0) * (np.sin(bx * (xx[m] - cx)) * np.cos(by * (yy[m] - cy)))
    hy[m] += by * (h1 - h0) * (np.cos(bx * (xx[m] - cx)) * np.sin(by * (yy[m] - cy)))
    return h, hx, hy


# ----------------------------------------------------------------------------
# integrate.py
# ----------------------------------------------------------------------------

def predictor_corrector(q, p, tau, direction):
    """integrate.py:38-77 with hyperbolicFlux (:133-165) and diffusiveFlux (:168-198)."""
    Fx = np.zeros_like(q)
    Fy = np.zeros_like(q)
    Fx[0] = q[1]
    Fx[1] = p
    Fy[0] = q[2]
    Fy[2] = p
    Dx = np.zeros_like(q)
    Dy = np.zeros_like(q)
    Dx[1], Dx[2] = tau[0], tau[2]
    Dy[1], Dy[2] = tau[2], tau[1]
    Fx = Fx + Dx
    Fy = Fy + Dy
    fx = -direction * (np.roll(Fx, direction, axis=1) - Fx)
    fy = -direction * (np.roll(Fy, direction, axis=2) - Fy)
    return fx, fy


def source(q, h, stress, lower, upper):
    """integrate.py:80-130."""
    out = np.zeros_like(q)
    out[0] = (-q[1] * h[1] - q[2] * h[2]) / h[0]
    out[1] = ((stress[0] - upper[0]) * h[1] + (stress[2] - upper[5]) * h[2] + upper[4] - lower[4]) / h[0]
    out[2] = ((stress[2] - upper[5]) * h[1] + (stress[1] - upper[1]) * h[2] + upper[3] - lower[3]) / h[0]
    return out


# ----------------------------------------------------------------------------
# problem.py driver
# ----------------------------------------------------------------------------

class FVOracle:
    """Restated ``Problem`` hot loop for fixed-form closures (no GP, no elastic wall).

    Parameters are the reference's sanitised dicts (io.py:139-395): ``grid`` with
    Nx, Ny, dx, dy, Lx, Ly and the 12 ``bc_*`` mask lists + ``bc_*_D_val``; ``geo``
    with U, V, type, ...; ``prop`` with shear, bulk, EOS + parameters; ``numerics``
    with dt, adaptive, CFL, MC_order, tol, max_it.
    """

    def __init__(self, grid, geo, prop, numerics, extra=None, topo=None, write_freq=1000):
        self.grid, self.geo, self.prop, self.numerics = grid, geo, prop, numerics
        self.write_freq = write_freq
        shape = (grid['Nx'] + 2, grid['Ny'] + 2)
        # problem.py:662-670
        self.q = np.empty((3, *shape))
        self.q[0] = prop['rho0']
        self.q[1] = prop['rho0'] * geo['U'] / 2.0
        self.q[2] = prop['rho0'] * geo['V'] / 2.0
        self.topo = make_topography(grid, geo) if topo is None else np.array(topo, dtype=float)
        self.extra = np.zeros((1, *shape))          # problem.py:131-135
        if extra is not None:
            self.extra[...] = extra
        self.pressure = np.zeros(shape)
        self.wall_xz = np.zeros((12, *shape))
        self.wall_yz = np.zeros((12, *shape))
        self.bulk = np.zeros((3, *shape))
        self.kinetic_energy_old = self.kinetic_energy
        self._stop = False
        # problem.py:412-443 (_pre_run without the GP parts)
        self.step = 0
        self.simtime = 0.
        self.residual = 1.
        self.residual_buffer = deque([self.residual], 5)
        self.dt = numerics['CFL'] * self.dt_crit if numerics['adaptive'] else numerics['dt']
        self.tol = numerics['tol']
        self.max_it = numerics['max_it']

    # -- scalar properties, problem.py:314-362 (all over the padded arrays) ----
    @property
    def q_is_valid(self):
        return (not np.any(np.isnan(self.q))) and (not np.any(self.q[0] < 0.))

    @property
    def mass(self):
        return np.sum(self.q[0] * self.topo[0] * self.grid['dx'] * self.grid['dy'])

    @property
    def kinetic_energy(self):
        return np.sum((self.q[1]**2 + self.q[2]**2) / self.q[0] / 2.)

    @property
    def v_max(self):
        return np.sqrt((self.q[1]**2 + self.q[2]**2) / self.q[0]).max()

    @property
    def v_sound(self):
        return C.eos_sound_velocity(self.q[0], self.prop).max()     # stress.py:538-539

    @property
    def dt_crit(self):
        return min(self.grid['dx'], self.grid['dy']) / (self.v_max + self.v_sound)

    @property
    def cfl(self):
        return self.dt / self.dt_crit

    @property
    def converged(self):
        return bool(np.all(np.array(self.residual_buffer) < self.tol))

    # -- closures, stress.py --------------------------------------------------
    def closures(self):
        """Pressure.update (:600-622), WallStress.update x2 (:289-362), BulkStress.update
        (:427-459), fixed-form branch; viscosity modifiers via viscosity()."""
        g, pr = self.geo, self.prop
        h = self.topo[:3]
        Ls = self.extra[0]
        self.pressure[...] = C.eos_pressure(self.q[0], pr)
        eta = self.viscosity()
        sb = C.stress_bottom(self.q, h, g['U'], g['V'], eta, pr['bulk'], Ls)
        st = C.stress_top(self.q, h, g['U'], g['V'], eta, pr['bulk'], Ls)
        for f, k in ((self.wall_xz, 4), (self.wall_yz, 3)):
            f[:3] = sb[:3] / 2.
            f[6:9] = st[:3] / 2.
            f[5] = sb[-1] / 2.
            f[11] = st[-1] / 2.
            f[k] = sb[k]
            f[k + 6] = st[k]
        self.bulk[...] = C.stress_avg(self.q, h, g['U'], g['V'], eta, pr['bulk'], Ls)

    def viscosity(self):
        """Scalar, or the (Nx+2, Ny+2) field of stress.py:305-326 when a piezo / thinning block is present;
        needs self.pressure of the current q."""
        if 'piezo' not in self.prop and 'thinning' not in self.prop:
            return self.prop['shear']
        xx, yy = midpoint_grid(self.grid)
        return C.shear_viscosity(self.pressure, self.q[0], self.topo[0], xx[:, 0], yy[0, :],
                                 self.geo['U'], self.geo['V'], self.prop)

    # -- ghost cells, problem.py:676-768 ---------------------------------------
    def ghost_values(self, bc_type, axis, direction):
        g, q = self.grid, self.q
        if axis == 0:
            side = 'xE' if direction > 0 else 'xW'
            mask = g[f'bc_{side}_{bc_type}']
            adj = q[mask, -2:-1, :] if direction > 0 else q[mask, 1:2, :]
        else:
            side = 'yS' if direction > 0 else 'yN'
            mask = g[f'bc_{side}_{bc_type}']
            adj = q[mask, :, -2:-1] if direction > 0 else q[mask, :, 1:2]
        target = g[f'bc_{side}_D_val']      # read for 'N' too (problem.py:738-753): KeyError when that side has no 'D' (io.py:183-191)
        a1, a2, q2 = 0.5, 0.0, 0.0
        if bc_type == 'D':
            return (target - a1 * adj + a2 * q2) / (a1 - a2)
        return ((1.0 - a1) * adj + a2 * q2) / (a1 - a2)

    def fill_ghosts(self):
        g, q = self.grid, self.q
        if all(g['bc_xE_P']):
            q[:, 0, :] = q[:, -2, :].copy()
        else:
            q[g['bc_xE_D'], :1, :] = self.ghost_values('D', 0, -1)
            q[g['bc_xE_N'], :1, :] = self.ghost_values('N', 0, -1)
        if all(g['bc_xW_P']):
            q[:, -1, :] = q[:, 1, :].copy()
        else:
            q[g['bc_xW_D'], -1:, :] = self.ghost_values('D', 0, 1)
            q[g['bc_xW_N'], -1:, :] = self.ghost_values('N', 0, 1)
        if all(g['bc_yS_P']):
            q[:, :, 0] = q[:, :, -2].copy()
        else:
            q[g['bc_yS_D'], :, :1] = self.ghost_values('D', 1, -1)
            q[g['bc_yS_N'], :, :1] = self.ghost_values('N', 1, -1)
        if all(g['bc_yN_P']):
            q[:, :, -1] = q[:, :, 1].copy()
        else:
            q[g['bc_yN_D'], :, -1:] = self.ghost_values('D', 1, 1)
            q[g['bc_yN_N'], :, -1:] = self.ghost_values('N', 1, 1)

    # -- one MacCormack step, problem.py:509-610 --------------------------------
    def directions(self):
        mc = self.numerics['MC_order']
        switch = (self.step % 2 == 0) * 2 - 1 if mc == 0 else mc
        return [[-1, 1], [1, -1]][(switch + 1) // 2]

    def one_pass(self, d):
        dx, dy, dt = self.grid['dx'], self.grid['dy'], self.dt
        self.closures()
        fX, fY = predictor_corrector(self.q, self.pressure, self.bulk, d)
        src = source(self.q, self.topo, self.bulk,
                     self.wall_xz[:6] + self.wall_yz[:6], self.wall_xz[6:] + self.wall_yz[6:])
        self.q[...] = self.q - dt * (fX / dx + fY / dy - src)
        self.fill_ghosts()

    def update(self):
        q0 = self.q.copy()
        for d in self.directions():
            self.one_pass(d)
        self.q[...] = (self.q + q0) / 2.0
        if self.q_is_valid:
            self._post_update()
        else:
            self._finalize(q0)

    def _post_update(self):
        self.fill_ghosts()
        ekin = self.kinetic_energy
        self.residual = abs(ekin - self.kinetic_energy_old) / self.kinetic_energy_old / self.cfl
        self.residual_buffer.append(self.residual)
        self.kinetic_energy_old = ekin
        self.step += 1
        self.simtime += self.dt
        if self.numerics['adaptive']:
            self.dt = self.numerics['CFL'] * self.dt_crit

    def _finalize(self, q0):
        self.q[...] = q0
        self.closures()
        self._stop = True

    def run(self, max_steps=None):
        n = 0
        while not self.converged and self.step < self.max_it and not self._stop:
            self.update()
            n += 1
            if max_steps is not None and n >= max_steps:
                break
        return n
