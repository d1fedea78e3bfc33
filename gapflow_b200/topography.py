"""Host-side gap geometry (runs once at set-up; its output is an INPUT of the hot path).

Same generators and midpoint grid as the reference (GaPFlow/topography.py:38-170, 227-255):
journal, inclined, parabolic, cdc, asperity(num=1); `flip` swaps the axes.  Returns the
reference's ``topography`` field: (4, Nx+2, Ny+2) = h, dh/dx, dh/dy, deformation (zero).
"""
import numpy as np


def midpoint_grid(grid, rows=None):
    """Cell midpoints of the padded grid; ``rows=(a, b)`` keeps padded x-rows a..b-1 only
    (a slab of the global grid: same element-wise arithmetic, so values are identical)."""
    def axis(n, L, sl=None):
        idx = np.arange(-1, n + 1)
        if sl is not None:
            idx = idx[sl[0]:sl[1]]
        return idx / n * L + (L / n) / 2.
    return np.meshgrid(axis(grid['Nx'], grid['Lx'], rows), axis(grid['Ny'], grid['Ly']), indexing='ij')


def _journal(xx, yy, grid, geo):
    w = 2. * np.pi / grid['Lx']
    if 'eps' in geo:
        mean = geo['CR'] / w
        amp = geo['eps'] * mean
    else:
        amp, mean = (geo['hmax'] - geo['hmin']) / 2., (geo['hmax'] + geo['hmin']) / 2.
    return mean + amp * np.cos(w * xx), -amp * w * np.sin(w * xx), np.zeros_like(xx)


def _inclined(xx, yy, grid, geo):
    s = (geo['hmin'] - geo['hmax']) / grid['Lx']
    return geo['hmax'] + s * xx, np.ones_like(xx) * s, np.zeros_like(xx)


def _parabolic(xx, yy, grid, geo):
    Lx = grid['Lx']
    a = 4. / Lx**2 * (geo['hmax'] - geo['hmin'])
    return a * (xx - Lx / 2.)**2 + geo['hmin'], 2 * a * (xx - Lx / 2.), np.zeros_like(xx)


def _cdc(xx, yy, grid, geo):
    Lx, lo, hi, b = grid['Lx'], geo['hmin'], geo['hmax'], geo['b']
    s = (hi - lo) / (Lx / 2 - 2 * b)
    zones = [((xx >= b) & (xx < Lx / 2 - b), lambda x: hi - s * (x - b), -s),
             ((xx >= Lx / 2 - b) & (xx < Lx / 2 + b), lambda x: lo + 0 * x, 0.),
             ((xx >= Lx / 2 + b) & (xx < Lx - b), lambda x: lo + s * (x - (Lx / 2 + b)), s)]
    h, hx = np.ones_like(xx) * hi, np.zeros_like(xx)
    for m, f, slope in zones:
        h[m] = f(xx[m])
        hx[m] = slope
    return h, hx, np.zeros_like(xx)


def _asperity(xx, yy, grid, geo):
    if geo.get('num', 1) != 1:
        raise NotImplementedError("asperity num > 1 draws unseeded random heights in the reference")
    lo, hi, Lx, Ly = geo['hmin'], geo['hmax'], grid['Lx'], grid['Ly']
    m = ((xx // Lx).astype(int) == 0) & ((yy // Ly).astype(int) == 0)
    kx, ky = np.pi / Lx, np.pi / Ly
    X, Y = xx[m] - np.mean(xx[m]), yy[m] - np.mean(yy[m])
    h, hx, hy = np.ones_like(xx) * hi, np.zeros_like(xx), np.zeros_like(xx)
    h[m] -= (hi - lo) * (np.cos(kx * X) * np.cos(ky * Y))
    hx[m] += kx * (hi - lo) * (np.sin(kx * X) * np.cos(ky * Y))
    hy[m] += ky * (hi - lo) * (np.cos(kx * X) * np.sin(ky * Y))
    return h, hx, hy


_GENERATORS = {'journal': _journal, 'inclined': _inclined, 'parabolic': _parabolic, 'cdc': _cdc,
               'asperity': _asperity}


def make_topography(grid, geo, rows=None):
    """``rows=(a, b)``: only padded x-rows a..b-1 of the global field (slab decomposition)."""
    if rows is not None and (geo['type'] == 'asperity' or geo.get('flip', False)):
        raise NotImplementedError("row-range generation needs a generator that is point-wise in x")
    xx, yy = midpoint_grid(grid, rows)
    h, hx, hy = _GENERATORS[geo['type']](xx, yy, grid, geo)
    out = np.zeros((4, *h.shape))
    if geo.get('flip', False):
        out[0], out[1], out[2] = h.T, hy.T, hx.T
    else:
        out[0], out[1], out[2] = h, hx, hy
    return out
