"""Derives the wall and gap-averaged viscous stresses of the lubrication film for Navier slip at the top wall only
("top") and at both walls ("both"), INCLUDING the in-plane gradient terms (dqx, dqy), from the physics -- not from the
reference's expanded formulas -- and writes gapflow_b200/csrc/closures_slip_gen.cuh.

Model (GaPFlow paper / models/profiles.py): across the gap the in-plane velocity is the parabola u(z) = a z^2 + b z + c
fixed by
    bottom wall (z = 0):  u = U            ("top")       u - Ls du/dz = U      ("both")
    top wall    (z = h):  u + Ls du/dz = 0
    flux:                 (1/h) int_0^h u dz = jx / rho
(v likewise with V, jy).  The viscous stress of a compressible Newtonian fluid with the lubrication ordering (w and its
derivatives dropped) is, in Voigt order xx, yy, zz, yz, xz, xy,
    t_xx = v1 u_x + v2 v_y,  t_yy = v2 u_x + v1 v_y,  t_zz = v2 (u_x + v_y),  t_yz = eta v_z,  t_xz = eta u_z,
    t_xy = eta (u_y + v_x),          v1 = zeta + 4/3 eta,  v2 = zeta - 2/3 eta,
where the in-plane derivatives are taken at fixed z through h(x, y), rho(x, y), jx(x, y), jy(x, y).
stress_bottom = the tensor at z = 0, stress_top at z = h, stress_avg = (1/h) int_0^h (xx, yy, xy) dz.

    python tools/derive_slip_closures.py            # derive, check against the reference where it is present, write the header
"""
import os
import sys

import sympy as sp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gapflow_b200", "csrc", "closures_slip_gen.cuh")

z, h, hx, hy, Ls, U, V, eta, zeta = sp.symbols('z h hx hy Ls U V eta zeta', real=True)
q0, q1, q2 = sp.symbols('q0 q1 q2', real=True)
dx0, dx1, dx2, dy0, dy1, dy2 = sp.symbols('dx0 dx1 dx2 dy0 dy1 dy2', real=True)
v1 = zeta + sp.Rational(4, 3) * eta
v2 = zeta - sp.Rational(2, 3) * eta


def profile(mode, wall, flux):
    a, b, c = sp.symbols('a b c')
    u = a * z**2 + b * z + c
    du = sp.diff(u, z)
    bottom = (u - (Ls * du if mode == 'both' else 0)).subs(z, 0) - wall
    top = (u + Ls * du).subs(z, h)
    mean = sp.integrate(u, (z, 0, h)) / h - flux / q0
    sol = sp.solve([bottom, top, mean], [a, b, c], dict=True)[0]
    return sp.simplify(u.subs(sol))


def tensors(mode):
    u = profile(mode, U, q1)
    v = profile(mode, V, q2)

    def d_dx(f):
        return sp.diff(f, h) * hx + sp.diff(f, q0) * dx0 + sp.diff(f, q1) * dx1 + sp.diff(f, q2) * dx2

    def d_dy(f):
        return sp.diff(f, h) * hy + sp.diff(f, q0) * dy0 + sp.diff(f, q1) * dy1 + sp.diff(f, q2) * dy2
    ux, uy, vx, vy = d_dx(u), d_dy(u), d_dx(v), d_dy(v)
    tau = [v1 * ux + v2 * vy, v2 * ux + v1 * vy, v2 * (ux + vy), eta * sp.diff(v, z), eta * sp.diff(u, z), eta * (uy + vx)]
    bot = [sp.simplify(t.subs(z, 0)) for t in tau]
    top = [sp.simplify(t.subs(z, h)) for t in tau]
    avg = [sp.simplify(sp.integrate(tau[k], (z, 0, h)) / h) for k in (0, 1, 5)]
    return bot, top, avg


def check_against_reference(derived):
    sys.path.insert(0, ROOT)
    from oracle import ref_loader
    if not ref_loader.available():
        print("reference not present: derivation not checked here")
        return
    import numpy as np
    R = ref_loader.modules()['viscous']
    rng = np.random.default_rng(0)
    syms = [h, hx, hy, Ls, U, V, eta, zeta, q0, q1, q2, dx0, dx1, dx2, dy0, dy1, dy2]
    worst = 0.0
    for mode, (bot, top, avg) in derived.items():
        fb, ft, fa = (sp.lambdify(syms, e, 'numpy') for e in (bot, top, avg))
        for trial in range(200):
            n = 64
            q = np.stack([1.0 + 0.2 * rng.random(n), rng.standard_normal(n), rng.standard_normal(n)])
            hh = np.stack([1.0 + rng.random(n), 0.1 * rng.standard_normal(n), 0.1 * rng.standard_normal(n)])
            dqx, dqy = 0.3 * rng.standard_normal((3, n)), 0.3 * rng.standard_normal((3, n))
            if trial % 3 == 0:
                dqx, dqy = np.zeros((3, n)), np.zeros((3, n))
            ls = rng.random(n) * (trial % 2)
            u_, v_, e_, z_ = rng.standard_normal(), rng.standard_normal(), 0.5 + rng.random(), rng.random()
            args = [hh[0], hh[1], hh[2], ls, u_, v_, e_, z_, q[0], q[1], q[2], dqx[0], dqx[1], dqx[2], dqy[0], dqy[1], dqy[2]]
            for name, f in (('stress_bottom', fb), ('stress_top', ft), ('stress_avg', fa)):
                want = getattr(R, name)(q, hh, u_, v_, e_, z_, ls, dqx=dqx, dqy=dqy, slip=mode)
                got = np.stack([np.broadcast_to(np.asarray(c, dtype=float), (n,)) for c in f(*args)])
                err = np.abs(got - want).max() / max(np.abs(want).max(), 1e-300)
                worst = max(worst, err)
                assert err < 1e-10, (mode, name, trial, err)
    print(f"derived closed forms == reference functions (slip top / both, with gradients): max rel err {worst:.2e}")


def emit(derived):
    from sympy.printing.c import C99CodePrinter

    class P(C99CodePrinter):
        def _print_Pow(self, e):               # small integer powers as products
            b, n = e.as_base_exp()
            if n.is_Integer and 2 <= int(n) <= 4:
                s = self._print(b)
                return "(" + "*".join(["(" + s + ")"] * int(n)) + ")"
            if n.is_Integer and -4 <= int(n) <= -1:
                s = self._print(b)
                return "(1.0/(" + "*".join(["(" + s + ")"] * (-int(n))) + "))"
            return super()._print_Pow(e)
    pr = P({'precision': 17})
    lines = ["// GENERATED by tools/derive_slip_closures.py -- do not edit.  Wall and gap-averaged viscous stresses of the",
             "// lubrication film with Navier slip at the top wall (\"top\") or at both walls (\"both\"), in-plane gradients of",
             "// (rho, jx, jy) included: the closed forms that GaPFlow/models/viscous.py:37-786 tabulates, re-derived from the",
             "// parabolic profile + Newtonian stress with sympy (common sub-expressions factored).  Tolerance contract: these",
             "// agree with the reference's expanded polynomials to rounding (<= 1e-12 of the largest component), not bit for bit;",
             "// the time step itself never calls them (it uses slip = \"top\", zero gradients: closures.cuh, bit-exact).",
             "#pragma once", "namespace gpf {", "struct SlipIn { double h, hx, hy, Ls, U, V, eta, zeta, q0, q1, q2, dx0, dx1, dx2, dy0, dy1, dy2; };", ""]
    for mode, (bot, top, avg) in derived.items():
        for name, exprs in (('bottom', bot), ('top', top), ('avg', avg)):
            repl, red = sp.cse(exprs, symbols=sp.numbered_symbols('t'), optimizations='basic')
            lines.append(f"__host__ __device__ inline void slip_{mode}_{name}(const SlipIn& in, double* out) {{")
            lines.append("    const double h = in.h, hx = in.hx, hy = in.hy, Ls = in.Ls, U = in.U, V = in.V, eta = in.eta, zeta = in.zeta;")
            lines.append("    const double q0 = in.q0, q1 = in.q1, q2 = in.q2, dx0 = in.dx0, dx1 = in.dx1, dx2 = in.dx2, dy0 = in.dy0, dy1 = in.dy1, dy2 = in.dy2;")
            lines.append("    (void)hx; (void)hy; (void)U; (void)V; (void)zeta; (void)dx0; (void)dx1; (void)dx2; (void)dy0; (void)dy1; (void)dy2; (void)q1; (void)q2;")
            for s_, e_ in repl:
                lines.append(f"    const double {s_} = {pr.doprint(e_)};")
            for k, e_ in enumerate(red):
                lines.append(f"    out[{k}] = {pr.doprint(e_)};")
            lines.append("}")
            lines.append("")
    lines.append("}  // namespace gpf")
    with open(OUT, "w") as f:
        f.write("\n".join(lines) + "\n")
    print("wrote", OUT, len(lines), "lines")


if __name__ == "__main__":
    derived = {mode: tensors(mode) for mode in ('top', 'both')}
    check_against_reference(derived)
    emit(derived)
