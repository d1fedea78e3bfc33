// Fused MacCormack pass kernel + its launcher, compiled twice (see fv_pass_exact.cu / fv_pass_fast.cu):
//   GPF_ARITH_NS = exact, -fmad=false : every product / sum rounded separately in the reference's order,
//                  IEEE quotients (Markstein)            -> bit-identical to the reference's NumPy path
//   GPF_ARITH_NS = fastm, -fmad=true, GPF_FAST_ARITH=1 : FMA contraction, quotients as a * rcp(b)
//                  (<= 2 ulp each)                        -> within 1e-12 relative per step (north_star)
#pragma once
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include "fv_pass.cuh"
#include "fv_host.h"

#ifndef GPF_ARITH_NS
#error "define GPF_ARITH_NS (exact | fastm) before including fv_pass_kernel.cuh"
#endif

namespace gpf {
namespace GPF_ARITH_NS {

// ---- shared-memory carve-up of the pass kernel (every TMA destination 128-byte aligned) -------
constexpr int BOX = BX * BY;                                    // doubles per staged plane
constexpr size_t al128(size_t x) { return (x + 127) / 128 * 128; }
constexpr size_t BOX3_BYTES = al128((size_t)3 * BOX * 8);
constexpr size_t BOX2_BYTES = al128((size_t)2 * BOX * 8);
constexpr size_t BOX1_BYTES = al128((size_t)BOX * 8);
constexpr size_t OWN3_BYTES = (size_t)3 * TX * TY * 8;          // q0 of the tile's own cells (corrector)

// NST: depth of the TMA ring (tiles in flight ahead of the one being computed: NST - 1)
template <bool CORR, bool LS, bool XF, int NST, bool VF = false>
struct Smem {
    static constexpr size_t Q = 0;
    static constexpr size_t TOPO = Q + BOX3_BYTES;
    static constexpr size_t LSO = TOPO + (XF ? BOX2_BYTES : BOX3_BYTES);    // x-driven film: dh/dy plane not loaded
    static constexpr size_t ETA = LSO + (LS ? BOX1_BYTES : 0);               // per-cell shear viscosity (piezo / thinning)
    static constexpr size_t Q0 = ETA + (VF ? BOX1_BYTES : 0);
    static constexpr size_t STAGE = Q0 + (CORR ? OWN3_BYTES : 0);
    static constexpr size_t CLOS = NST * STAGE;                 // p, t0, t1, t2 of the staged cells
    static constexpr size_t BAR = CLOS + al128((size_t)4 * BOX * 8);
    static constexpr size_t RED = BAR + 128;                    // <= 4 mbarriers, 2 tile flags
    static constexpr size_t TOTAL = RED + (FV_THREADS / 32) * 4 * 8;
    static constexpr uint32_t TX_BYTES = (uint32_t)(3 * BOX * 8) + (uint32_t)((XF ? 2 : 3) * BOX * 8) +
                                         (LS ? (uint32_t)(BOX * 8) : 0u) + (VF ? (uint32_t)(BOX * 8) : 0u) +
                                         (CORR ? (uint32_t)OWN3_BYTES : 0u);
};



// ---- per-cell arithmetic of the two phases (FAST: branch-free divisions, see closures.cuh) ----
// The straight-line quotients need every operand in [2^-100, 2^100] (jx, jy may also be exactly 0).
// Topography, slip length and the scalar parameters are checked once when they are uploaded
// (FvParams::static_fast); the solution is checked here, per staged cell, on the integer pipe.
__device__ __forceinline__ bool q_in_fast_range(const Cell& c) {
    return in_fast_range_nz(c.q0) & in_fast_range(c.q1) & in_fast_range(c.q2);
}

// what the straight-line (GEN = false) kernel must know about a staged cell: with contracted
// arithmetic only denominators matter (a * rcp(b): the density; gap height and parameters are
// static), the exact contract also needs the numerators of its Markstein quotients in range
__device__ __forceinline__ bool cell_in_fast_range(const Cell& c) {
#if GPF_FAST_ARITH
    return in_fast_range_nz(c.q0);
#else
    return q_in_fast_range(c);
#endif
}

struct Clos4 { double p, t0, t1, t2; };

// viscosity scalars of one cell: the configuration's, or (VF) those of the cell's own shear viscosity
// eta (stress.py:305-326 makes `eta` a field; viscous.py:82-83 then evaluates v1, v2 per cell)
template <bool VF>
__device__ __forceinline__ Visc cell_visc(const FvParams& P, double eta) {
    Visc m = reinterpret_cast<const Visc&>(P.visc);
    if constexpr (VF) {
        m.eta = eta;
        m.v1 = m.zeta + 4. / 3. * eta;
        m.v2 = m.zeta - 2. / 3. * eta;
        m.two_eta = 2 * eta;
        m.two_v2 = 2 * m.v2;
    }
    return m;
}

template <bool LS, int EOS, bool FAST, bool XF, bool VF = false>
__device__ __forceinline__ Clos4 phase1_cell(const Cell& c, const FvParams& P, double eta = 0.0) {
    Clos4 r;
#if GPF_FAST_ARITH
    if constexpr (FAST && !LS) {
        // zero slip length: the gap-averaged stress (viscous.py:663-715) reduces to
        //   (xx, yy, xy) = (v1 A1 + v2 A2, v2 A1 + v1 A2, eta (hx jy + hy jx) a),  a = 1/(h rho),
        //   A1 = hx jx a, A2 = hy jy a   (same closed form, common factor h rho cancelled)
        const Visc& m = reinterpret_cast<const Visc&>(P.visc);
        r.p = eos_pressure<EOS, true>(c.q0, P.eos);
        const double a = rcp_exact(c.h * c.q0);
        const double A1 = c.hx * c.q1 * a;
        if constexpr (XF) {
            r.t0 = m.v1 * A1;
            r.t1 = m.v2 * A1;
            r.t2 = m.eta * (c.hx * c.q2 * a);
        } else {
            const double A2 = c.hy * c.q2 * a;
            r.t0 = m.v1 * A1 + m.v2 * A2;
            r.t1 = m.v2 * A1 + m.v1 * A2;
            r.t2 = m.eta * ((c.hx * c.q2 + c.hy * c.q1) * a);
        }
        return r;
    }
#endif
    r.p = eos_pressure<EOS, FAST>(c.q0, P.eos);
    if constexpr (VF) {
        const Visc m = cell_visc<true>(P, eta);
        stress_avg<LS, FAST, XF>(c, m, r.t0, r.t1, r.t2);
    } else {
        stress_avg<LS, FAST, XF>(c, reinterpret_cast<const Visc&>(P.visc), r.t0, r.t1, r.t2);
    }
    return r;
}
template <bool LS, int EOS, bool VF = false>
__device__ __noinline__ Clos4 phase1_cell_slow(const Cell c, const FvParams& P, double eta = 0.0) {
    return phase1_cell<LS, EOS, false, false, VF>(c, P, eta);
}

struct Nb { double q1x, Ax, t2x, q2y, t2y, Cy; };   // neighbour fluxes: Fx = (q1, A, t2), Fy = (q2, t2, C)
struct Q3 { double v0, v1, v2; };
// GP surrogate (SG variant): predictive means that replace the own shear component of the wall stresses
// (stress.py:350-356: field[_out_index] = mean[0], field[_out_index + 6] = mean[1]); mask bit 0: xz model,
// bit 1: yz model
struct SgWall { double lo_xz, up_xz, lo_yz, up_yz; unsigned mask; };

// fluxes (integrate.py:65-75), wall stresses (stress.py:328-362), source (integrate.py:118-128)
// and the update q - dt*(fX/dx + fY/dy - src) (problem.py:558) of one interior cell.
// flux = -d*(F[nb] - F[own]): the sign flip is exact and done on the integer pipe.  The xz + yz
// field sums of problem.py:554-555 are u/2 + u/2 == u (exact) and u + 0.
template <bool LS, bool FAST, bool XF, bool VF = false, bool SG = false>
__device__ __forceinline__ Q3 phase2_cell(const Cell& c, const Clos4& k, const Nb& n, const FvParams& P,
                                          unsigned sbit, double dt, double eta = 0.0, const SgWall* sg = nullptr) {
    static_assert(!SG || !FAST, "the surrogate variant runs the plain-division path");
    static_assert(!VF || LS, "the viscosity-field variants use the general (slip-length) closures");
    Visc Vcell;
    if constexpr (VF) Vcell = cell_visc<true>(P, eta);
    const Visc& V = VF ? Vcell : reinterpret_cast<const Visc&>(P.visc);
    const double fx0 = flip(n.q1x - c.q1, sbit);
    const double fx1 = flip(n.Ax - (k.p + k.t0), sbit);
    const double fx2 = flip(n.t2x - k.t2, sbit);
    const double fy0 = flip(n.q2y - c.q2, sbit);
    const double fy1 = flip(n.t2y - k.t2, sbit);
    const double fy2 = flip(n.Cy - (k.p + k.t1), sbit);
#if GPF_FAST_ARITH
    if constexpr (FAST && !LS) {
        // zero slip length: with a = 1/(h rho), yh = 1/h, e1 = (3 jx - U rho) a, e2 = (3 jy - V rho) a the
        // wall stresses (viscous.py:88-105, 333-426) are
        //   top:    xx = 2 (v1 hx e1 + v2 hy e2), yy = 2 (v2 hx e1 + v1 hy e2), xy = 2 eta (hx e2 + hy e1),
        //           xz = -2 eta e1, yz = -2 eta e2;      bottom: xz = 2 eta (e1 - U yh), yz = 2 eta (e2 - V yh)
        // so  up4 - lo4 = 2 eta (U yh - 2 e1),  up3 - lo3 = 2 eta (V yh - 2 e2)   (integrate.py:118-128)
        const double yh = rcp_exact(c.h);
        const double a = rcp_exact(c.h * c.q0);
        const double e1 = (3.0 * c.q1 - V.U * c.q0) * a;
        const double w4 = V.two_eta * (V.U * yh - 2.0 * e1);
        double s0, s1, s2;
        if constexpr (XF) {
            const double e2 = 3.0 * c.q2 * a;
            const double u0 = 2.0 * (V.v1 * c.hx * e1);
            const double u5 = V.two_eta * (c.hx * e2);
            const double w3 = V.two_eta * (-2.0 * e2);
            s0 = -(c.q1 * c.hx) * yh;
            s1 = ((k.t0 - u0) * c.hx + w4) * yh;
            s2 = ((k.t2 - u5) * c.hx + w3) * yh;
        } else {
            const double e2 = (3.0 * c.q2 - V.V * c.q0) * a;
            const double u0 = 2.0 * (V.v1 * c.hx * e1 + V.v2 * c.hy * e2);
            const double u1 = 2.0 * (V.v2 * c.hx * e1 + V.v1 * c.hy * e2);
            const double u5 = V.two_eta * (c.hx * e2 + c.hy * e1);
            const double w3 = V.two_eta * (V.V * yh - 2.0 * e2);
            s0 = -(c.q1 * c.hx + c.q2 * c.hy) * yh;
            s1 = ((k.t0 - u0) * c.hx + (k.t2 - u5) * c.hy + w4) * yh;
            s2 = ((k.t2 - u5) * c.hx + (k.t1 - u1) * c.hy + w3) * yh;
        }
        Q3 q;
        q.v0 = c.q0 - dt * (fx0 * P.rdx + fy0 * P.rdy - s0);
        q.v1 = c.q1 - dt * (fx1 * P.rdx + fy1 * P.rdy - s1);
        q.v2 = c.q2 - dt * (fx2 * P.rdx + fy2 * P.rdy - s2);
        return q;
    }
#endif
    double b3, b4, u[6];
    stress_bottom<LS, FAST, XF>(c, V, b3, b4);
    stress_top<LS, false, FAST, XF>(c, V, u);
    if constexpr (SG) {
        if (sg->mask & 1u) { b4 = sg->lo_xz; u[4] = sg->up_xz; }
        if (sg->mask & 2u) { b3 = sg->lo_yz; u[3] = sg->up_yz; }
    }
    const double rh = rcp_sel<FAST>(c.h);
    double s0, s1, s2;
    if constexpr (XF) {      // hy == 0: the hy terms are exact zeros
        s0 = div_sel<FAST>(-c.q1 * c.hx, c.h, rh);
        s1 = div_sel<FAST>((k.t0 - u[0]) * c.hx + u[4] - b4, c.h, rh);
        s2 = div_sel<FAST>((k.t2 - u[5]) * c.hx + u[3] - b3, c.h, rh);
    } else {
        s0 = div_sel<FAST>(-c.q1 * c.hx - c.q2 * c.hy, c.h, rh);
        s1 = div_sel<FAST>((k.t0 - u[0]) * c.hx + (k.t2 - u[5]) * c.hy + u[4] - b4, c.h, rh);
        s2 = div_sel<FAST>((k.t2 - u[5]) * c.hx + (k.t1 - u[1]) * c.hy + u[3] - b3, c.h, rh);
    }
    Q3 q;
    q.v0 = c.q0 - dt * (div_sel<FAST>(fx0, P.dx, P.rdx) + div_sel<FAST>(fy0, P.dy, P.rdy) - s0);
    q.v1 = c.q1 - dt * (div_sel<FAST>(fx1, P.dx, P.rdx) + div_sel<FAST>(fy1, P.dy, P.rdy) - s1);
    q.v2 = c.q2 - dt * (div_sel<FAST>(fx2, P.dx, P.rdx) + div_sel<FAST>(fy2, P.dy, P.rdy) - s2);
    return q;
}
template <bool LS, bool VF = false>
__device__ __noinline__ Q3 phase2_cell_slow(const Cell c, const Clos4 k, const Nb n, const FvParams& P,
                                            unsigned sbit, double dt, double eta = 0.0) {
    return phase2_cell<LS, false, false, VF>(c, k, n, P, sbit, dt, eta);
}
template <bool LS>
__device__ __noinline__ Q3 phase2_cell_sg(const Cell c, const Clos4 k, const Nb n, const FvParams& P,
                                          unsigned sbit, double dt, const SgWall sg) {
    return phase2_cell<LS, false, false, false, true>(c, k, n, P, sbit, dt, 0.0, &sg);
}

// exact flags of one stored cell, off the hot path (problem.py:320-332)
__device__ __noinline__ int cell_flags_slow(double a, double b, double c, double w) {
    int f = 0;
    if (a != a || b != b || c != c) f |= FLAG_NAN;
    if (a < 0.0) f |= FLAG_NEG;
    if (!(w >= 0.0)) f |= FLAG_VNAN;
    return f;
}

// ---- fused slab exchange: what the general launch of a pass does for its rank (one CTA, all threads) -------------
// Raises the neighbours' flags for this pass's edge rows (stored remotely by the boundary tiles, see
// boundary_cell) and -- corrector -- folds this rank's per-CTA records in a fixed order and stores the result, with
// its sequence number, into every rank's record table: the jobs of fv_halo_send_kernel and fv_record_send_kernel
// without their launches.  `scratch`: >= 5 * FV_THREADS doubles of shared memory.
__device__ __forceinline__ void slab_publish(const PassArgs& A, bool corr, double* scratch) {
    const SlabX& X = A.X;
    const int tid = threadIdx.x;
    if (tid == 0) {
        __threadfence_system();                      // the edge rows (this kernel's or the straight-line launch's) first
        const unsigned long long seq = *X.xseq + 1;
        if (X.peer_flag[0]) st_release_sys(X.peer_flag[0], seq);
        if (X.peer_flag[1]) st_release_sys(X.peer_flag[1], seq);
        *X.xseq = seq;
    }
    if (!corr) return;
    double* sE = scratch; double* sV = sE + FV_THREADS; double* sC = sV + FV_THREADS; double* sM = sC + FV_THREADS;
    int* sF = reinterpret_cast<int*>(sM + FV_THREADS);
    double E = 0.0, VM = -INFINITY, CM = -INFINITY, M = 0.0;
    int F = 0;
    for (int b = tid; b < X.nrec; b += FV_THREADS) {
        E += __ldcg(&A.partial[b].ekin);
        M += __ldcg(&A.partial[b].mass);
        VM = fmax(VM, __ldcg(&A.partial[b].vmax2));
        CM = fmax(CM, __ldcg(&A.partial[b].cs2));
        F |= __ldcg(&A.pflags[b]);
    }
    sE[tid] = E; sV[tid] = VM; sC[tid] = CM; sM[tid] = M; sF[tid] = F;
    __syncthreads();
    for (int st = FV_THREADS / 2; st > 0; st >>= 1) {
        if (tid < st) {
            sE[tid] += sE[tid + st];
            sM[tid] += sM[tid + st];
            sV[tid] = fmax(sV[tid], sV[tid + st]);
            sC[tid] = fmax(sC[tid], sC[tid + st]);
            sF[tid] |= sF[tid + st];
        }
        __syncthreads();
    }
    if (tid < X.world) {
        const unsigned long long seq = *X.rseq + 1;            // this fold; fv_finish_kernel advances *rseq
        const int parity = (int)(seq & 1ull);
        double* dst = reinterpret_cast<double*>(X.peers[tid] + X.off_recs) + ((size_t)parity * X.world + X.rank) * 8;
        const int cf = *reinterpret_cast<const volatile int*>(X.err) ? FLAG_COMM : 0;
        dst[0] = sE[0]; dst[1] = sV[0]; dst[2] = sC[0]; dst[3] = sM[0]; dst[4] = (double)(sF[0] | cf);
        __threadfence_system();
        st_release_sys(reinterpret_cast<unsigned long long*>(X.peers[tid] + X.off_rflag) + X.rank, seq);
    }
}

// Persistent kernel: fv_ctas_per_sm(CORR) CTAs per SM walk the tiles round-robin.  A 2-deep ring of
// TMA-staged boxes keeps the loads of tile n+1 in flight under the arithmetic of tile n.
//
// Every pass is two launches of this template:
//   GEN = false  straight-line arithmetic only (reciprocal-based quotients valid for operands in
//                [2^-100, 2^100]).  It range-checks what it divides by on the integer pipe and, if
//                any cell (or dt, or a static field) is outside, raises A.redo[CORR]; its output is
//                then garbage.  No out-of-line call sits in its loop: the general kernel's cold calls
//                cost 20 % of the pass through the registers they pin (r1h -> r1i).
//   GEN = true   the general kernel (tile-uniform choice between the same straight-line code and
//                plain IEEE divisions, exact validity flags).  Returns at once unless A.redo[CORR] is
//                raised, else recomputes the whole pass: a pass reads only buffers it does not write.
// Each GEN = false launch lowers the other pass's flag; a stale raised flag only costs time.
//
// SG = true (GP surrogate closures, gpf_fv_set_surrogate): only the general kernel is launched and every
// tile takes the plain-division path; the pressure of a staged cell and the own shear component of the
// wall stresses come from the GP means of the pass's input (dense arrays over the padded grid, row
// i*(Ny+2)+j like gp.py:223-232 orders them) times the model's Yscale (gp.py:362).
// SX = true: the fused slab exchange (SlabX).  A separate instantiation: folded into the plain kernels as a run-time
// switch it cost them 10 registers and 1.3 % (contracted) to 5 % (exact arithmetic) of a single-GPU step (r2g).
template <bool CORR, bool LS, int EOS, bool XF, bool GEN, bool VF = false, bool SG = false, bool SX = false>
__global__ void __launch_bounds__(FV_THREADS, fv_ctas_per_sm(CORR)) fv_pass_kernel(const __grid_constant__ PassArgs A) {
    static_assert(!SG || (GEN && LS && !XF && !VF && EOS == -1), "surrogate variant: general kernel, general closures");
    static_assert(!SX || (!VF && !SG), "fused exchange: fixed-form variants only");
    constexpr int NST = fv_nstage(CORR);
    using S = Smem<CORR, LS, XF, NST, VF>;
    extern __shared__ __align__(1024) unsigned char smem[];      // no static shared memory: offset 0
    uint64_t* const bar = reinterpret_cast<uint64_t*>(smem + S::BAR);
    double* const sp = reinterpret_cast<double*>(smem + S::CLOS);
    double* const st0 = sp + BOX;
    double* const st1 = st0 + BOX;
    double* const st2 = st1 + BOX;
    int* const sbad = reinterpret_cast<int*>(smem + S::BAR + 32);
    double* const sred = reinterpret_cast<double*>(smem + S::RED);

    const FvParams& P = A.P;
    const DevState* st = A.st;
    // run-loop condition (problem.py:401) evaluated on the device; uniform per grid
    // fused slab exchange: the general launch closes the pass for its rank (flags, record) also when it has nothing
    // to recompute, and when the run has halted (every rank halts in the same step; the fold still runs)
    if (st->stop || (A.until_converged && st->halt)) {
        if (GEN && SX && blockIdx.x == 0) slab_publish(A, CORR, reinterpret_cast<double*>(smem));
        return;
    }
    if (GEN && !SG && A.redo[CORR] == 0) {
        if (SX && blockIdx.x == 0) slab_publish(A, CORR, reinterpret_cast<double*>(smem));
        return;
    }

    const int tid = threadIdx.x;
    const int cur = st->cur;
    const int in_idx = CORR ? (cur + 1) % 3 : cur;
    const int out_idx = CORR ? (cur + 2) % 3 : (cur + 1) % 3;
    // MacCormack order (problem.py:521-522)
    const int sw = (P.mc_order == 0) ? (((st->step & 1LL) == 0) ? 1 : -1) : P.mc_order;
    const int d = CORR ? -sw : sw;
    const double dt = st->dt;
    const int ox = (d > 0) ? 1 : 0;         // own cells sit at box offset (ox, oy)
    constexpr int oy = 2;                   // box origin = own column - 2 (even -> 16-byte aligned)
    const int Nx = P.L.Nx, Ny = P.L.Ny;
    const size_t plane = P.L.plane;
    const int pitch = P.L.pitch;
    const int nty = P.nblk_x;               // tiles along y

    // tiles are walked round-robin: tile t = (bx, by), bx = t / nty along x, by = t % nty along y.
    // The walk advances (bx, by) by a fixed step, so no division runs in the loop.
    const int step_bx = (int)gridDim.x / nty, step_by = (int)gridDim.x % nty;

    // a launch covers all tile rows (mul 1, off 0), the interior ones (mul 1, off 1) or the first
    // and the last one (mul = rows - 1, off 0): see FvSolver::pass
    const int bx_mul = A.bx_mul, bx_off = A.bx_off;
    // Fused slab exchange: a pass with d > 0 needs the lower neighbour's edge row in its FIRST tile row, one with d < 0
    // the upper neighbour's in its LAST.  The rows arrive when the neighbour's previous pass ends, so the tile rows are
    // walked such that the one that needs them comes last (d > 0: top to bottom reversed): every wait then has a
    // whole pass of slack and ranks may drift apart by up to a pass without stalling each other.
    const bool flip_rows = SX && d > 0;
    auto tile_row = [&](int vbx) { const int r = vbx * bx_mul + bx_off; return flip_rows ? P.nblk_y - 1 - r : r; };
    // fused slab exchange: the ghost row a first / last tile row stages was stored by the neighbour's previous pass;
    // its flag carries the number of that exchange = the exchanges this rank has published so far
    const unsigned long long xneed = (SX && tid == 0) ? *A.X.xseq : 0ull;
    auto issue = [&](int vbx, int by, int stage) {         // thread 0 only
        if constexpr (SX) {
            const int trow = tile_row(vbx);
            bool waited = false;
            if (ox == 1 && trow == 0 && P.bc.type[0][0] == GPF_BC_HALO) { spin_until(A.X.my_flag + 0, xneed, A.X.err); waited = true; }
            if (ox == 0 && trow == P.nblk_y - 1 && P.bc.type[1][0] == GPF_BC_HALO) { spin_until(A.X.my_flag + 1, xneed, A.X.err); waited = true; }
            if (waited) asm volatile("fence.proxy.async;" ::: "memory");     // peer-written rows, read next by the TMA unit
        }
        const int i0 = 1 + tile_row(vbx) * TX, j0 = 1 + by * TY;
        unsigned char* base = smem + (size_t)stage * S::STAGE;
        uint64_t* b = bar + stage;
        mbar_expect_tx(b, S::TX_BYTES);
        const int cm = (j0 - oy) + 1;       // memory column of the box origin (j0 = 1 + 64k -> even)
        const int cx = i0 - ox;
        tma_load_3d(base + S::Q, &A.qmap[in_idx], b, cm, cx, 0);
        tma_load_3d(base + S::TOPO, XF ? &A.topo2_map : &A.topo_map, b, cm, cx, 0);
        if (LS) tma_load_3d(base + S::LSO, &A.ls_map, b, cm, cx, 0);
        if (VF) tma_load_3d(base + S::ETA, &A.eta_map, b, cm, cx, 0);
        if (CORR) tma_load_3d(base + S::Q0, &A.q0map[cur], b, j0 + 1, i0, 0);
    };

    if (tid == 0) {
        for (int k = 0; k < NST; ++k) mbar_init(bar + k, 1);
        fence_mbar_init();
        sbad[0] = sbad[1] = 0;
    }
    __syncthreads();
    int tile = blockIdx.x;
    int bx = tile / nty, by = tile % nty;
    // issue cursor: runs NST - 1 tiles ahead of the tile being computed
    int itile = tile, ibx = bx, iby = by;
    auto advance = [&](int& t, int& x, int& y) {
        t += (int)gridDim.x;
        x += step_bx;
        y += step_by;
        if (y >= nty) { y -= nty; ++x; }
    };
#pragma unroll
    for (int k = 0; k < NST - 1; ++k) {
        if (tid == 0 && itile < A.ntiles) issue(ibx, iby, k);
        advance(itile, ibx, iby);
    }

    // thread -> cells: column ty, rows 2g and 2g+1 of the tile
    const int ty = tid & (TY - 1);
    const int g = tid >> 6;
    static_assert(FV_THREADS / TY * 2 == TX && TY == 64, "thread-to-cell map assumes 2 rows per thread");
    const int o_own = (2 * g + ox) * BY + ty + oy;                 // box offset of the first own cell
    // halo cell of this thread (warps 0-1: the halo row; warp 2: the halo column, lanes >= TX repeat
    // cell TX-1 so the warp stays convergent; the duplicate stores write identical values)
    const int hrow = (d > 0) ? 0 : TX, hcol = (d > 0) ? oy - 1 : oy + TY;
    const int hcl = (tid - TY < TX - 1) ? tid - TY : TX - 1;
    const int o_halo = (tid < TY) ? hrow * BY + tid + oy : (hcl + ox) * BY + hcol;
    const bool has_halo = tid < TY + 32;                           // warp-uniform
    const int on_dx = -d * BY, on_dy = -d;                        // offsets to the neighbours differenced against
    const unsigned sbit = (d > 0) ? 0x80000000u : 0u;             // -d * x == x with its sign bit flipped when d = +1
    const bool dt_fast = in_fast_range_nz(dt) && P.static_fast;
    if (!GEN) {
        if (tid == 0 && blockIdx.x == 0) A.redo[CORR ? 0 : 1] = 0;
        if (!dt_fast) {                                            // grid-uniform
            if (tid == 0) A.redo[CORR] = 1;
            return;
        }
    }
    bool bad = false;                                              // GEN = false: an operand left the fast range

    double* const qout = A.qbuf[out_idx];
    const double* const q0b = A.qbuf[cur];
    // this thread's first own cell of tile (0, 0); a tile adds bx*TX rows and by*TY columns
    const size_t a_thread = (size_t)(1 + 2 * g) * pitch + (1 + ty) + 1;
    Red red;
    red.ekin = 0.0;
    red.vmax2 = -INFINITY;
    red.cs2 = -INFINITY;
    red.flags = 0;
    constexpr bool DEFER_C = CORR && (EOS == GPF_EOS_DH);
    double run_thr = -INFINITY;                                    // (1 - 2^-40) x the largest density this warp has seen

    auto store_clos = [&](int o, const Clos4& k) { sp[o] = k.p; st0[o] = k.t0; st1[o] = k.t1; st2[o] = k.t2; };
    int stage = 0;
    unsigned phase = 0;
    for (int it = 0; tile < A.ntiles; ++it) {
        const int fl = it & 1;                                     // this tile's slow-path flag
        // the stage read by tile it-1 was released by that tile's closing CTA barrier: refill it
        {
            const int is = (stage == 0) ? NST - 1 : stage - 1;
            if (tid == 0 && itile < A.ntiles) issue(ibx, iby, is);
            advance(itile, ibx, iby);
        }
        const unsigned char* base = smem + (size_t)stage * S::STAGE;
        const double* sq = reinterpret_cast<const double*>(base + S::Q);
        const double* stp = reinterpret_cast<const double*>(base + S::TOPO);
        const double* sls = reinterpret_cast<const double*>(base + S::LSO);
        const double* sq0 = reinterpret_cast<const double*>(base + S::Q0);
        const double* seta = reinterpret_cast<const double*>(base + S::ETA);
        mbar_wait(bar + stage, phase);
        // the cell's shear viscosity (VF) must be a divisor-safe operand like the other static fields
        auto load_eta = [&](int o) { return VF ? seta[o] : 0.0; };
        auto eta_ok = [&](double e) { return !VF || in_fast_range_nz(e); };

        auto load_cell = [&](int o) {
            Cell c;
            c.q0 = sq[o];
            c.q1 = sq[BOX + o];
            c.q2 = sq[2 * BOX + o];
            c.h = stp[o];
            c.hx = stp[BOX + o];
            c.hy = XF ? 0.0 : stp[2 * BOX + o];
            c.Ls = LS ? sls[o] : 0.0;
            return c;
        };

        // ---- phase 1: pressure + gap-averaged stress on tile + halo row + halo column ----
        if constexpr (!GEN) {
            const Cell ca = load_cell(o_own), cb = load_cell(o_own + BY);
            const double ea = load_eta(o_own), eb = load_eta(o_own + BY);
            bad |= !(cell_in_fast_range(ca) & cell_in_fast_range(cb) & eta_ok(ea) & eta_ok(eb));
            store_clos(o_own, phase1_cell<LS, EOS, true, XF, VF>(ca, P, ea));
            store_clos(o_own + BY, phase1_cell<LS, EOS, true, XF, VF>(cb, P, eb));
            if (has_halo) {
                const Cell ch = load_cell(o_halo);
                const double eh = load_eta(o_halo);
                bad |= !(cell_in_fast_range(ch) & eta_ok(eh));
                store_clos(o_halo, phase1_cell<LS, EOS, true, XF, VF>(ch, P, eh));
            }
        } else {
            const Cell ca = load_cell(o_own), cb = load_cell(o_own + BY);
            const double ea = load_eta(o_own), eb = load_eta(o_own + BY);
            const bool ok = !SG & q_in_fast_range(ca) & q_in_fast_range(cb) & eta_ok(ea) & eta_ok(eb);
            if (__all_sync(0xffffffffu, ok) && P.static_fast) {   // warp-uniform
                store_clos(o_own, phase1_cell<LS, EOS, true, XF, VF>(ca, P, ea));
                store_clos(o_own + BY, phase1_cell<LS, EOS, true, XF, VF>(cb, P, eb));
            } else {
                store_clos(o_own, phase1_cell_slow<LS, EOS, VF>(ca, P, ea));
                store_clos(o_own + BY, phase1_cell_slow<LS, EOS, VF>(cb, P, eb));
                sbad[fl] = 1;             // this tile takes the plain-division path in phase 2
            }
        }
        if (GEN && has_halo) {
            const Cell ch = load_cell(o_halo);
            const double eh = load_eta(o_halo);
            if (__all_sync(0xffffffffu, !SG & q_in_fast_range(ch) & eta_ok(eh)) && P.static_fast) {
                store_clos(o_halo, phase1_cell<LS, EOS, true, XF, VF>(ch, P, eh));
            } else {
                store_clos(o_halo, phase1_cell_slow<LS, EOS, VF>(ch, P, eh));
                sbad[fl] = 1;
            }
        }
        if constexpr (SG) {
            // the GP pressure of every staged cell that lies inside the padded array (the others are
            // neighbours of cells that are neither stored nor counted)
            if (A.sg_p) {
                const int io = 1 + tile_row(bx) * TX - ox, jo = 1 + by * TY - oy;   // cell of box offset 0
                auto put = [&](int o) {
                    const int r = o / BY, i = io + r, j = jo + (o - r * BY);
                    if (i >= 0 && i <= Nx + 1 && j >= 0 && j <= Ny + 1)
                        sp[o] = A.sg_p[(size_t)i * (Ny + 2) + j] * A.sg_scale[0];
                };
                put(o_own);
                put(o_own + BY);
                if (has_halo) put(o_halo);
            }
        }
        __syncthreads();

        // ---- phase 2: wall stresses, fluxes, source, update, ghosts, reductions ----------
        const bool fast = !SG && (!GEN || ((sbad[fl] == 0) && dt_fast));           // tile-uniform
        if (GEN && tid == 0) sbad[fl ^ 1] = 0;                                      // next tile's flag
        const int rbx = tile_row(bx);                                              // tile row in the grid
        const int i0 = 1 + rbx * TX, j0 = 1 + by * TY;
        const bool full = (i0 + TX - 1 <= Nx) && (j0 + TY - 1 <= Ny);
        const bool btile = (rbx == 0) | (i0 + TX - 1 >= Nx) | (by == 0) | (j0 + TY - 1 >= Ny);
        const size_t a_tile = a_thread + (size_t)(rbx * TX) * pitch + (size_t)(by * TY);
        double rho[2] = {-INFINITY, -INFINITY};
        auto update_cell = [&](int k) {
            const int o = o_own + k * BY;
            const Cell c = load_cell(o);
            Clos4 kk;
            kk.p = sp[o]; kk.t0 = st0[o]; kk.t1 = st1[o]; kk.t2 = st2[o];
            Nb nb;
            nb.q1x = sq[BOX + o + on_dx]; nb.Ax = sp[o + on_dx] + st0[o + on_dx]; nb.t2x = st2[o + on_dx];
            nb.q2y = sq[2 * BOX + o + on_dy]; nb.t2y = st2[o + on_dy]; nb.Cy = sp[o + on_dy] + st1[o + on_dy];
            const double e = load_eta(o);
            if constexpr (SG) {
                SgWall w;
                w.mask = (A.sg_xz ? 1u : 0u) | (A.sg_yz ? 2u : 0u);
                w.lo_xz = w.up_xz = w.lo_yz = w.up_yz = 0.0;
                const int i = 1 + tile_row(bx) * TX + 2 * g + k, j = 1 + by * TY + ty;
                if (i <= Nx && j <= Ny) {
                    const size_t np = (size_t)(Nx + 2) * (Ny + 2), a = (size_t)i * (Ny + 2) + j;
                    if (A.sg_xz) { w.lo_xz = A.sg_xz[a] * A.sg_scale[1]; w.up_xz = A.sg_xz[np + a] * A.sg_scale[1]; }
                    if (A.sg_yz) { w.lo_yz = A.sg_yz[a] * A.sg_scale[2]; w.up_yz = A.sg_yz[np + a] * A.sg_scale[2]; }
                }
                return phase2_cell_sg<LS>(c, kk, nb, P, sbit, dt, w);
            } else
            if constexpr (GEN) {
                return fast ? phase2_cell<LS, true, XF, VF>(c, kk, nb, P, sbit, dt, e)
                            : phase2_cell_slow<LS, VF>(c, kk, nb, P, sbit, dt, e);
            } else {
                return phase2_cell<LS, true, XF, VF>(c, kk, nb, P, sbit, dt, e);
            }
        };
        // corrector (116 registers at 2 CTAs/SM): both cells of a thread in one straight-line block,
        // two independent FP64 chains; the predictor's 80-register budget has no room for that
        constexpr bool ILP2 = !GEN && CORR;
        Q3 q2c[2];
        if constexpr (ILP2) {
            q2c[0] = update_cell(0);
            q2c[1] = update_cell(1);
        }
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            // cells of a ragged tile that lie outside the grid are computed (on zero-filled or
            // foreign data) but neither stored nor counted: every lane must reach the warp votes
            const bool act = full || ((i0 + 2 * g + k <= Nx) && (j0 + ty <= Ny));
            const Q3 qq = ILP2 ? q2c[k] : update_cell(k);
            double qn0 = qq.v0, qn1 = qq.v1, qn2 = qq.v2;
            double w = 0.0;                                        // (jx^2 + jy^2) / rho of the stored cell (corrector)
            if (CORR) {
                const int oo = (2 * g + k) * TY + ty;
                qn0 = (qn0 + sq0[oo]) * 0.5;                       // problem.py:563 ((q + q0)/2.0, exact)
                qn1 = (qn1 + sq0[TX * TY + oo]) * 0.5;
                qn2 = (qn2 + sq0[2 * TX * TY + oo]) * 0.5;
                // kinetic energy and speed (problem.py:342, 347): w = (jx^2 + jy^2)/rho
                const double num = qn1 * qn1 + qn2 * qn2;
                if constexpr (GEN) {
                    const bool wok = fast & in_fast_range_nz(qn0) & in_fast_range(num);
                    if (__all_sync(0xffffffffu, wok)) w = div_sel<true>(num, qn0, rcp_exact(qn0));
                    else w = num / qn0;
                } else {
                    // an invalid cell (NaN, negative density) is never in range: the general kernel
                    // then redoes the pass and sets the exact flags
                    w = div_sel<true>(num, qn0, rcp_exact(qn0));
#if GPF_FAST_ARITH
                    if (act) bad |= !in_fast_range_nz(qn0) | !(w >= 0.0) | (__double2hiint(qn0) < 0);
#else
                    if (act) bad |= !(in_fast_range_nz(qn0) & in_fast_range(num)) | !(w >= 0.0) | (__double2hiint(qn0) < 0);
#endif
                }
                if (act) {
                    red.ekin += w * 0.5;
                    red.vmax2 = fmax(red.vmax2, w);
                    // NaN anywhere in the cell makes w NaN; a negative density has its sign bit set
                    if (GEN && (!(w >= 0.0) || __double2hiint(qn0) < 0)) red.flags |= cell_flags_slow(qn0, qn1, qn2, w);
                    if (DEFER_C) rho[k] = qn0;
                    else red_accum_c<EOS>(red, qn0, P.eos);
                }
            }
            if (act) {
                double* dst = qout + a_tile + (size_t)k * pitch;
                dst[0] = qn0;
                dst[plane] = qn1;
                dst[2 * plane] = qn2;
            }
            // ghost cells derived from this cell (problem.py:560 / 576)
            if (btile) {
                const int i = i0 + 2 * g + k, j = j0 + ty;
                if (act && ((i == 1) | (i == Nx) | (j == 1) | (j == Ny))) {
                    Red br;
                    if constexpr (SX) br = boundary_cell<CORR, EOS, true>(P, i, j, qq.v0, qq.v1, qq.v2, qn0, qn1, qn2, qout, q0b, &A.X, out_idx);
                    else br = boundary_cell<CORR, EOS>(P, i, j, qq.v0, qq.v1, qq.v2, qn0, qn1, qn2, qout, q0b);
                    if (CORR) {
                        // periodic ring, fused exchange: this row is also the global ghost row of the reference's padded
                        // array (it lives in the neighbour's memory): the cell and its y-ghost images count once more
                        if (SX && ((A.X.edge_dup[0] && i == 1) || (A.X.edge_dup[1] && i == Nx))) red.ekin += w * 0.5 + br.ekin;
                        red.ekin += br.ekin;
                        red.vmax2 = fmax(red.vmax2, br.vmax2);
                        red.cs2 = fmax(red.cs2, br.cs2);
                        red.flags |= br.flags;
                    }
                }
            }
        }
        if (DEFER_C) {
            // v_sound = max sqrt(dp/drho).  For Dowson-Higginson dp/drho = K/(C2*rho0 - rho)^2 rises
            // monotonically with rho below the pole, by >= 7e-13 relative per 2^-40 relative step in
            // rho (for rho >= rho0/2), while its 7-operation FP64 evaluation errs by < 1e-15.  So the
            // cell that attains the maximum of the ROUNDED values lies within 2^-40 of the largest
            // density: only cells that close to the largest density this warp has met so far are
            // evaluated -- bit-identical to evaluating every cell.  Outside the proven range every
            // cell is evaluated.  Positive doubles order like their bit patterns: the warp's exact
            // maximum takes two 32-bit REDUX operations.
            const double m = fmax(rho[0], rho[1]);
            const int whi = __reduce_max_sync(0xffffffffu, __double2hiint(m));
            const unsigned wlo = __reduce_max_sync(0xffffffffu, (__double2hiint(m) == whi) ? (unsigned)__double2loint(m) : 0u);
            const double wmax = __hiloint2double(whi, (int)wlo);
            const bool ok = (wmax >= 0.5 * P.eos.p[0]) && (wmax <= P.eos.p[4]);
            if (!ok) {
#pragma unroll
                for (int k = 0; k < 2; ++k)
                    if (rho[k] != -INFINITY) red_accum_c<EOS>(red, rho[k], P.eos);
            } else if (wmax >= run_thr) {
                const double thr = wmax * (1.0 - 0x1p-40);
#pragma unroll
                for (int k = 0; k < 2; ++k)
                    if (rho[k] >= thr) red_accum_c<EOS>(red, rho[k], P.eos);
                run_thr = fmax(run_thr, thr);
            }
        }
        __syncthreads();          // closures and this stage's boxes are free again
        advance(tile, bx, by);
        if (++stage == NST) { stage = 0; phase ^= 1u; }
    }

    if (!GEN && bad) A.redo[CORR] = 1;
    if (CORR) {
        // warp shuffles, then one record per CTA (fixed order: deterministic)
        double e = warp_sum(red.ekin);
        double vm = warp_max(red.vmax2);
        double cm = warp_max(red.cs2);
        int fl = red.flags;
#pragma unroll
        for (int oo = 16; oo > 0; oo >>= 1) fl |= __shfl_xor_sync(0xffffffffu, fl, oo);
        const int w = tid >> 5, lane = tid & 31;
        if (lane == 0) {
            sred[w * 4 + 0] = e;
            sred[w * 4 + 1] = vm;
            sred[w * 4 + 2] = cm;
            sred[w * 4 + 3] = (double)fl;
        }
        __syncthreads();
        if (tid == 0) {
            double E = 0.0, VM = -INFINITY, CM = -INFINITY;
            int F = 0;
            for (int ww = 0; ww < FV_THREADS / 32; ++ww) {
                E += sred[ww * 4 + 0];
                VM = fmax(VM, sred[ww * 4 + 1]);
                CM = fmax(CM, sred[ww * 4 + 2]);
                F |= (int)sred[ww * 4 + 3];
            }
            const int slot = A.rec0 + (int)blockIdx.x;
            A.partial[slot].ekin = E;
            A.partial[slot].vmax2 = VM;
            A.partial[slot].cs2 = CM;
            A.partial[slot].mass = 0.0;
            A.pflags[slot] = F;
        }
    }
    if (GEN && SX) {
        // the general kernel ran the pass in full (redo, or the surrogate variant): the last CTA out closes it
        __syncthreads();
        if (tid == 0) {
            __threadfence_system();                                // this CTA's record and (remote) edge rows before its exit count
            sbad[0] = (atomicAdd(A.X.xdone, 1u) == gridDim.x - 1) ? 1 : 0;
            if (sbad[0]) { *A.X.xdone = 0u; __threadfence(); }
        }
        __syncthreads();
        if (sbad[0]) slab_publish(A, CORR, reinterpret_cast<double*>(smem));
    }
}

template <bool CORR, bool LS, int EOS, bool XF, bool VF = false, bool SX = false>
static int launch_pass_variant(const PassArgs& A, int ncta, cudaStream_t s) {
    using Sm = Smem<CORR, LS, XF, fv_nstage(CORR), VF>;
    // the opt-in to > 48 KB of dynamic shared memory is a per-device attribute of the function
    static bool attr_set[64] = {};
    int dev = 0;
    GPF_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        GPF_CUDA(cudaFuncSetAttribute(fv_pass_kernel<CORR, LS, EOS, XF, false, VF, false, SX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Sm::TOTAL));
        GPF_CUDA(cudaFuncSetAttribute(fv_pass_kernel<CORR, LS, EOS, XF, true, VF, false, SX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Sm::TOTAL));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    fv_pass_kernel<CORR, LS, EOS, XF, false, VF, false, SX><<<ncta, FV_THREADS, Sm::TOTAL, s>>>(A);
    fv_pass_kernel<CORR, LS, EOS, XF, true, VF, false, SX><<<ncta, FV_THREADS, Sm::TOTAL, s>>>(A);
    return GPF_OK;
}

// surrogate variant: the general kernel alone
template <bool CORR>
static int launch_pass_surrogate(const PassArgs& A, int ncta, cudaStream_t s) {
    using Sm = Smem<CORR, true, false, fv_nstage(CORR), false>;
    static bool attr_set[64] = {};
    int dev = 0;
    GPF_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        GPF_CUDA(cudaFuncSetAttribute(fv_pass_kernel<CORR, true, -1, false, true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Sm::TOTAL));
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    fv_pass_kernel<CORR, true, -1, false, true, false, true><<<ncta, FV_THREADS, Sm::TOTAL, s>>>(A);
    return GPF_OK;
}

// variant dispatch: slip-length field, Dowson-Higginson inlined, x-driven film, viscosity field
template <bool CORR>
int launch_pass_dispatch(const PassArgs& A, bool ls, bool dh, bool xf, bool vf, int n, cudaStream_t s) {
    if (A.sg_p || A.sg_xz || A.sg_yz) return launch_pass_surrogate<CORR>(A, n, s);
    // viscosity field (piezo / thinning): one variant, the general closures with the slip-length plane
    // (zeros when the configuration has none: the LS forms with Ls = 0 round like the reference's)
    if (vf) return launch_pass_variant<CORR, true, -1, false, true>(A, n, s);
    if (A.X.on) {          // fused slab exchange: the SX instantiations
        switch ((ls ? 4 : 0) + (dh ? 2 : 0) + (xf ? 1 : 0)) {
        case 0: return launch_pass_variant<CORR, false, -1, false, false, true>(A, n, s);
        case 1: return launch_pass_variant<CORR, false, -1, true, false, true>(A, n, s);
        case 2: return launch_pass_variant<CORR, false, GPF_EOS_DH, false, false, true>(A, n, s);
        case 3: return launch_pass_variant<CORR, false, GPF_EOS_DH, true, false, true>(A, n, s);
        case 4: return launch_pass_variant<CORR, true, -1, false, false, true>(A, n, s);
        case 5: return launch_pass_variant<CORR, true, -1, true, false, true>(A, n, s);
        case 6: return launch_pass_variant<CORR, true, GPF_EOS_DH, false, false, true>(A, n, s);
        default: return launch_pass_variant<CORR, true, GPF_EOS_DH, true, false, true>(A, n, s);
        }
    }
    switch ((ls ? 4 : 0) + (dh ? 2 : 0) + (xf ? 1 : 0)) {
    case 0: return launch_pass_variant<CORR, false, -1, false>(A, n, s);
    case 1: return launch_pass_variant<CORR, false, -1, true>(A, n, s);
    case 2: return launch_pass_variant<CORR, false, GPF_EOS_DH, false>(A, n, s);
    case 3: return launch_pass_variant<CORR, false, GPF_EOS_DH, true>(A, n, s);
    case 4: return launch_pass_variant<CORR, true, -1, false>(A, n, s);
    case 5: return launch_pass_variant<CORR, true, -1, true>(A, n, s);
    case 6: return launch_pass_variant<CORR, true, GPF_EOS_DH, false>(A, n, s);
    default: return launch_pass_variant<CORR, true, GPF_EOS_DH, true>(A, n, s);
    }
}

}  // namespace GPF_ARITH_NS
}  // namespace gpf
