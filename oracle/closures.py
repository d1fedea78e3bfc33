"""NumPy restatement of GaPFlow's fixed-form closures -- TEST INFRASTRUCTURE ONLY.

Restates, for the only mode the driver uses (slip="top", dqx=dqy=0):

* EOS pressure          GaPFlow/models/pressure.py:35-325
* EOS sound speed       GaPFlow/models/sound.py:35-329
* wall / bulk stresses  GaPFlow/models/viscous.py:37-105 (bottom), :281-426 (top),
                        :612-715 (gap average)

Every product is evaluated strictly left to right and every sum in the reference's
term order, so the results are bit-identical to the reference's NumPy expressions
(terms that carry a zero ``dqx``/``dqy`` factor add an exact zero and are dropped).
``tests/test_oracle_vs_reference.py`` pins this against the reference's own
functions; ``tests/golden/fv_closures.npz`` carries the vectors to the GPU box.
"""
import os
import numpy as np

R_GAS = 8.31446261815324  # scipy.constants.gas_constant (exact SI value k_B*N_A), pressure.py:32

_BWR = None


def bwr_coeffs():
    """32 BWR coefficients (values of GaPFlow/models/bwr_coeffs.txt)."""
    global _BWR
    if _BWR is None:
        _BWR = np.loadtxt(os.path.join(os.path.dirname(__file__), "bwr_coeffs.txt"))
    return _BWR


def _prod(*f):
    """Left-to-right product, i.e. Python's ``a * b * c * ...``."""
    out = f[0]
    for x in f[1:]:
        out = out * x
    return out


# ----------------------------------------------------------------------------
# EOS pressure (pressure.py)
# ----------------------------------------------------------------------------

def eos_pressure(dens, prop):
    """pressure.py:35-76 dispatcher."""
    e = prop['EOS']
    if e == 'DH':      # pressure.py:79-109
        rho0, P0, C1, C2 = prop['rho0'], prop['P0'], prop['C1'], prop['C2']
        rho = np.minimum(dens, 0.99 * C2 * rho0)
        return P0 + (C1 * (rho / rho0 - 1.)) / (C2 - rho / rho0)
    if e == 'PL':      # pressure.py:112-137
        return prop['P0'] * (dens / prop['rho0'])**(1. / (1. - 0.5 * prop['alpha']))
    if e == 'vdW':     # pressure.py:140-173
        md = dens / prop['M'] * 1000.
        a = prop['a'] / 10.
        b = prop['b'] / 1000.
        return R_GAS * prop['T'] * md / (1. - b * md) - a * md**2
    if e == 'MT':      # pressure.py:176-205
        return prop['K'] / prop['n'] * ((dens / prop['rho0'])**prop['n'] - 1) + prop['P0']
    if e == 'cubic':   # pressure.py:208-230
        return prop['a'] * dens**3 + prop['b'] * dens**2 + prop['c'] * dens + prop['d']
    if e == 'BWR':     # pressure.py:233-274
        return _bwr_pressure(dens, prop['T'], prop.get('gamma', 3.))
    if e == 'Bayada':  # pressure.py:277-325
        return _bayada_pressure(dens, prop['rho_l'], prop['rho_v'], prop['c_l'], prop['c_v'])
    raise ValueError(e)


def _bwr_pressure(d, T, g):
    x = bwr_coeffs()
    poly = d * T
    poly = poly + d**2 * (x[0] * T + x[1] * np.sqrt(T) + x[2] + x[3] / T + x[4] / T**2)
    poly = poly + d**3 * (x[5] * T + x[6] + x[7] / T + x[8] / T**2)
    poly = poly + d**4 * (x[9] * T + x[10] + x[11] / T)
    poly = poly + d**5 * x[12]
    poly = poly + d**6 * (x[13] / T + x[14] / T**2)
    poly = poly + d**7 * (x[15] / T)
    poly = poly + d**8 * (x[16] / T + x[17] / T**2)
    poly = poly + d**9 * (x[18] / T**2)
    e = d**3 * (x[19] / T**2 + x[20] / T**3)
    e = e + d**5 * (x[21] / T**2 + x[22] / T**4)
    e = e + d**7 * (x[23] / T**2 + x[24] / T**3)
    e = e + d**9 * (x[25] / T**2 + x[26] / T**4)
    e = e + d**11 * (x[27] / T**2 + x[28] / T**3)
    e = e + d**13 * (x[29] / T**2 + x[30] / T**3 + x[31] / T**4)
    return poly + np.exp(-g * d**2) * e


def bayada_constants(rho_l, rho_v, c_l, c_v):
    N = rho_v * c_v**2 * rho_l * c_l**2 * (rho_v - rho_l) / (rho_v**2 * c_v**2 - rho_l**2 * c_l**2)
    Pcav = rho_v * c_v**2 - N * np.log(rho_v**2 * c_v**2 / (rho_l**2 * c_l**2))
    return N, Pcav


def _bayada_pressure(dens, rho_l, rho_v, c_l, c_v):
    N, Pcav = bayada_constants(rho_l, rho_v, c_l, c_v)
    dens = np.asarray(dens, dtype=float)
    alpha = (dens - rho_l) / (rho_v - rho_l)
    p = c_v**2 * dens
    lo = alpha < 0
    mix = np.logical_and(alpha <= 1, alpha >= 0)
    p[lo] = Pcav + (dens[lo] - rho_l) * c_l**2
    am = alpha[mix]
    den = rho_l * (rho_v * c_v**2 * (1 - am) + rho_l * c_l**2 * am)
    p[mix] = Pcav + N * np.log(rho_v * c_v**2 * dens[mix] / den)
    return p


# ----------------------------------------------------------------------------
# EOS sound speed (sound.py)
# ----------------------------------------------------------------------------

def eos_sound_velocity(dens, prop):
    """sound.py:35-81 dispatcher; returns sqrt(dp/drho) per cell."""
    e = prop['EOS']
    if e == 'DH':      # sound.py:84-111 (no density clamp here)
        rho0, C1, C2 = prop['rho0'], prop['C1'], prop['C2']
        g = C1 * rho0 * (C2 - 1.0) * (1 / dens) ** 2 / ((C2 * rho0 / dens - 1.0) ** 2)
        return np.sqrt(g)
    if e == 'PL':      # sound.py:114-139
        P0, rho0, al = prop['P0'], prop['rho0'], prop['alpha']
        g = -2.0 * P0 * (dens / rho0) ** (-2.0 / (al - 2.0)) / ((al - 2) * dens)
        return np.sqrt(g)
    if e == 'vdW':     # sound.py:142-174
        md = dens / prop['M'] * 1000.
        a = prop['a'] / 10.
        b = prop['b'] / 1000.
        g = R_GAS * prop['T'] / (1. - b * md)**2 - 2. * a * md
        return np.sqrt(g)
    if e == 'MT':      # sound.py:177-204
        g = prop['K'] / prop['rho0']**prop['n'] * dens ** (prop['n'] - 1)
        return np.sqrt(g)
    if e == 'cubic':   # sound.py:207-230
        g = 3 * prop['a'] * dens**2 + 2 * prop['b'] * dens + prop['c']
        return np.sqrt(g)
    if e == 'BWR':     # sound.py:233-293
        return np.sqrt(_bwr_dpdrho(dens, prop['T'], prop.get('gamma', 3.0)))
    if e == 'Bayada':  # sound.py:296-329
        rho_l, rho_v, c_l, c_v = prop['rho_l'], prop['rho_v'], prop['c_l'], prop['c_v']
        dens = np.asarray(dens, dtype=float)
        alpha = (dens - rho_l) / (rho_v - rho_l)
        mix = np.logical_and(alpha <= 1, alpha >= 0)
        c2 = np.ones_like(dens) * c_v**2
        c2[alpha < 0] = c_l**2
        c2[mix] = rho_v * rho_l * (c_v * c_l) ** 2 / (alpha[mix] * rho_l * c_l ** 2
                                                      + (1 - alpha[mix]) * rho_v * c_v**2) / dens[mix]
        return np.sqrt(c2)
    raise ValueError(e)


def _bwr_dpdrho(r, T, g):
    x = bwr_coeffs()
    c3 = (x[19] / T**2 + x[20] / T**3)
    c5 = (x[21] / T**2 + x[22] / T**4)
    c7 = (x[23] / T**2 + x[24] / T**3)
    c9 = (x[25] / T**2 + x[26] / T**4)
    c11 = (x[27] / T**2 + x[28] / T**3)
    c13 = (x[29] / T**2 + x[30] / T**3 + x[31] / T**4)
    ep = r**3 * c3 + r**5 * c5 + r**7 * c7 + r**9 * c9 + r**11 * c11 + r**13 * c13
    dep = (3.0 * r**2 * c3 + 5.0 * r**4 * c5 + 7.0 * r**6 * c7 + 9.0 * r**8 * c9
           + 11.0 * r**10 * c11 + 13.0 * r**12 * c13)
    out = T + 2.0 * r * (x[0] * T + x[1] * np.sqrt(T) + x[2] + x[3] / T + x[4] / T**2)
    out = out + 3.0 * r**2 * (x[5] * T + x[6] + x[7] / T + x[8] / T**2)
    out = out + 4.0 * r**3 * (x[9] * T + x[10] + x[11] / T)
    out = out + 5.0 * r**4 * x[12]
    out = out + 6.0 * r**5 * (x[13] / T + x[14] / T**2)
    out = out + 7.0 * r**6 * (x[15] / T)
    out = out + 8.0 * r**7 * (x[16] / T + x[17] / T**2)
    out = out + 9.0 * r**8 * (x[18] / T**2)
    out = out + np.exp(-g * r**2) * dep
    out = out - 2.0 * r * g * np.exp(-g * r**2) * ep
    return out


# ----------------------------------------------------------------------------
# Viscous stresses, slip="top", dqx = dqy = 0 (viscous.py)
# ----------------------------------------------------------------------------
# q = (rho, jx, jy); h = (gap, dh/dx, dh/dy); Voigt order xx, yy, zz, yz, xz, xy.
# Each stress component is  prefactor * sum(sign * prod(factors)) / denominator  with
# the factors multiplied left to right as the reference writes them.

def _sum_terms(terms):
    """terms: list of (sign, factors).  The reference's sums start with terms that
    are exactly zero (dq factors), so the first kept term enters as 0 +/- t = +/- t."""
    acc = None
    for sign, fac in terms:
        t = _prod(*fac)
        if acc is None:
            acc = t if sign > 0 else -t
        else:
            acc = acc + t if sign > 0 else acc - t
    return acc


def stress_bottom(q, h, U, V, eta, zeta, Ls):
    """Bottom-wall stress, viscous.py:88-105 (slip top): only yz and xz are non-zero."""
    q0, q1, q2 = q[0], q[1], q[2]
    h0 = h[0]
    tau = np.zeros((6, *q0.shape))
    den = h0 * q0 * (4 * Ls + h0)
    # note: the reference's first term is the product (-6) * Ls * V * q0
    tau[3] = 2 * eta * (-6 * Ls * V * q0 + 6 * Ls * q2 - 2 * V * h0 * q0 + 3 * h0 * q2) / den
    tau[4] = 2 * eta * (-6 * Ls * U * q0 + 6 * Ls * q1 - 2 * U * h0 * q0 + 3 * h0 * q1) / den
    return tau


def stress_top(q, h, U, V, eta, zeta, Ls):
    """Top-wall stress, viscous.py:333-426 (slip top) with dq = 0."""
    q0, q1, q2 = q[0], q[1], q[2]
    h0, hx, hy = h[0], h[1], h[2]
    v1 = zeta + 4 / 3 * eta
    v2 = zeta - 2 / 3 * eta
    q0s = q0 ** 2
    tau = np.zeros((6, *q0.shape))
    den = q0s * (16 * Ls**2 + 8 * Ls * h0 + h0 ** 2)

    def normal(va, vb):
        # viscous.py:335-355 (va=v1, vb=v2) and :357-378 (va=v2, vb=v1)
        return [(-1, (3, Ls, U, hx, q0s, va)), (-1, (3, Ls, V, hy, q0s, vb)),
                (+1, (9, Ls, hx, q0, q1, va)), (+1, (9, Ls, hy, q0, q2, vb)),
                (-1, (U, h0, hx, q0s, va)), (-1, (V, h0, hy, q0s, vb)),
                (+1, (3, h0, hx, q0, q1, va)), (+1, (3, h0, hy, q0, q2, vb))]

    tau[0] = 2 * _sum_terms(normal(v1, v2)) / den
    tau[1] = 2 * _sum_terms(normal(v2, v1)) / den
    # viscous.py:379-401: 2 * v2 * (...) / den, terms without the v factors
    zz = [(-1, (3, Ls, U, hx, q0s)), (-1, (3, Ls, V, hy, q0s)),
          (+1, (9, Ls, hx, q0, q1)), (+1, (9, Ls, hy, q0, q2)),
          (-1, (U, h0, hx, q0s)), (-1, (V, h0, hy, q0s)),
          (+1, (3, h0, hx, q0, q1)), (+1, (3, h0, hy, q0, q2))]
    tau[2] = 2 * v2 * _sum_terms(zz) / den
    tau[3] = 2 * eta * (V * q0 - 3 * q2) / (q0 * (4 * Ls + h0))   # viscous.py:402
    tau[4] = 2 * eta * (U * q0 - 3 * q1) / (q0 * (4 * Ls + h0))   # viscous.py:403
    # viscous.py:404-426
    xy = [(-1, (3, Ls, U, hy, q0s)), (-1, (3, Ls, V, hx, q0s)),
          (+1, (9, Ls, hx, q0, q2)), (+1, (9, Ls, hy, q0, q1)),
          (-1, (U, h0, hy, q0s)), (-1, (V, h0, hx, q0s)),
          (+1, (3, h0, hx, q0, q2)), (+1, (3, h0, hy, q0, q1))]
    tau[5] = 2 * eta * _sum_terms(xy) / den
    return tau


def stress_avg(q, h, U, V, eta, zeta, Ls):
    """Gap-averaged stress (xx, yy, xy), viscous.py:663-715 (slip top) with dq = 0."""
    q0, q1, q2 = q[0], q[1], q[2]
    h0, hx, hy = h[0], h[1], h[2]
    v1 = zeta + 4 / 3 * eta
    v2 = zeta - 2 / 3 * eta
    q0s = q0 ** 2
    tau = np.zeros((3, *q0.shape))
    den = h0 * q0s * (4 * Ls + h0)

    def normal(va, vb):
        return [(+1, (2, Ls, U, hx, q0s, va)), (+1, (2, Ls, V, hy, q0s, vb)),
                (-1, (2, Ls, hx, q0, q1, va)), (-1, (2, Ls, hy, q0, q2, vb)),
                (+1, (h0, hx, q0, q1, va)), (+1, (h0, hy, q0, q2, vb))]

    tau[0] = _sum_terms(normal(v1, v2)) / den
    tau[1] = _sum_terms(normal(v2, v1)) / den
    xy = [(+1, (2, Ls, U, hy, q0s)), (+1, (2, Ls, V, hx, q0s)),
          (-1, (2, Ls, hx, q0, q2)), (-1, (2, Ls, hy, q0, q1)),
          (+1, (h0, hx, q0, q2)), (+1, (h0, hy, q0, q1))]
    tau[2] = eta * _sum_terms(xy) / den
    return tau


# ---------------------------------------------------------------------------------
# Viscosity modifiers (row F11; CUDA: fv_viscosity_kernel in gapflow_b200/csrc/fv_kernels.cu)
# ---------------------------------------------------------------------------------
def piezoviscosity(p, mu0, piezo):
    """GaPFlow/models/viscosity.py:34-66 -> 144-259.  ``p`` is the pressure field, or the density for
    the Bayada-Chupin mixture laws (stress.py:306-309)."""
    name = piezo.get('name', 'none')
    if name == 'Barus':                  # viscosity.py:165
        return mu0 * np.exp(piezo['aB'] * p)
    if name == 'Roelands':               # viscosity.py:197
        return mu0 * np.exp(np.log(mu0 / piezo['mu_inf']) * (-1 + (1 + p / piezo['p_ref'])**piezo['z']))
    if name in ('Dukler', 'McAdams'):
        alpha = (p - piezo['rho_l']) / (piezo['rho_v'] - piezo['rho_l'])
        if name == 'Dukler':             # viscosity.py:227-228
            return alpha * piezo['eta_v'] + (1 - alpha) * mu0
        M = alpha * piezo['rho_v'] / p   # viscosity.py:257-259
        return piezo['eta_v'] * mu0 / (mu0 * M + piezo['eta_v'] * (1 - M))
    return np.ones_like(p) * mu0         # viscosity.py:64


def shear_rate_avg(dp_dx, dp_dy, h, u1, u2, mu):
    """viscosity.py:99-141 (Newtonian profile, mean of the absolute wall shear rates)."""
    grad_p = np.hypot(dp_dx, dp_dy)
    du_pois = h * grad_p / (2 * mu)
    du_carr = (u2 - u1) / h
    return (np.abs(du_pois + du_carr) + np.abs(-du_pois + du_carr)) / 2.


def shear_thinning_factor(shear_rate, mu0, thinning):
    """viscosity.py:69-96 -> 262-318: factor that multiplies the (piezo-)viscosity."""
    name = thinning.get('name', 'none')
    if name == 'Eyring':                 # viscosity.py:284-285
        tau0 = mu0 * shear_rate
        return thinning['tauE'] / tau0 * np.arcsinh(tau0 / thinning['tauE'])
    if name == 'Carreau':                # viscosity.py:316-318
        mu = thinning['mu_inf'] + (mu0 - thinning['mu_inf']) * \
            (1 + (thinning['lam'] * shear_rate)**thinning['a'])**((thinning['N'] - 1) / thinning['a'])
        return mu / mu0
    return np.ones_like(shear_rate)      # viscosity.py:94


def shear_viscosity(pressure, dens, h, x, y, U, V, prop):
    """The viscosity field both stress models build (stress.py:305-326 == 429-450): piezoviscosity of the
    CURRENT pressure field (density for the Bayada-Chupin EOS), then the thinning factor at the average
    shear rate, whose pressure gradient is np.gradient over the padded array with the cell-centre
    coordinates (stress.py:171-192: second-order central, one-sided at the array edges).  Note the
    reference passes (geo['U'], geo['V']) as (u1, u2) of shear_rate_avg (stress.py:317-318)."""
    mu0 = prop['shear']
    if 'piezo' in prop:
        mu0 = piezoviscosity(pressure if prop['EOS'] != 'Bayada' else dens, prop['shear'], prop['piezo'])
    if 'thinning' in prop:
        dp_dx = np.gradient(pressure, x, axis=0)
        dp_dy = np.gradient(pressure, y, axis=1)
        rate = shear_rate_avg(dp_dx, dp_dy, h, U, V, mu0)
        return mu0 * shear_thinning_factor(rate, mu0, prop['thinning'])
    return mu0
