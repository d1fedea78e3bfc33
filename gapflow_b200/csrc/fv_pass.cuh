// Fused MacCormack pass kernel (predictor or corrector) for sm_100a.
//
// One launch replaces, for every interior cell, the reference's per-pass chain
//   Pressure.update / WallStress.update x2 / BulkStress.update   (stress.py:289-362,427-459,600-622)
//   predictor_corrector + source                                   (integrate.py:38-130)
//   q = q - dt*(fX/dx + fY/dy - src)                               (problem.py:558)
//   _communicate_ghost_buffers                                     (problem.py:676-768)
// and, in the corrector, q = (q + q0)/2 (problem.py:563), the validity test
// (problem.py:320-332) and the reductions behind kinetic_energy, v_max and v_sound
// (problem.py:339-347, stress.py:538-539).
//
// Data path: one CTA per TX x TY tile.  Thread 0 issues TMA (cp.async.bulk.tensor.3d)
// loads of the q, topography and slip-length boxes (tile + one halo row and column on the
// side the pass differences towards) into shared memory and all threads wait on an
// mbarrier.  Phase 1 evaluates pressure and the gap-averaged stress once per staged cell
// into shared memory; phase 2 evaluates wall stresses, fluxes and source per interior
// cell, updates q and stores it coalesced.  Closures never touch HBM.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include "closures.cuh"

namespace gpf {

// ------------------------------ PTX helpers --------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t phase) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(phase)
        : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1,
                                            int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], "
        "[%2];" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}

// x with its sign bit XOR-ed (exact negation when sbit = 0x80000000), integer pipe
__device__ __forceinline__ double flip(double x, unsigned sbit) {
    return __hiloint2double(__double2hiint(x) ^ (int)sbit, __double2loint(x));
}

// ------------------------------ system-scope flags (slab exchange over peer memory) --------
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// Waits until *flag >= seq.  `err` is the communicator's error block: err[0] the error flag, the 64-bit word at
// err + 2 the wait bound in ns (gpf_comm_set_timeout).  A wait that exceeds the bound (a peer died) raises the
// flag instead of hanging the GPU; once it is up every later wait falls through at once, the records carry
// FLAG_COMM to every rank and fv_finish_kernel stops the run (DevState::comm_err).
__device__ __forceinline__ void spin_until(const unsigned long long* flag, unsigned long long seq, int* err) {
    if (ld_acquire_sys(flag) >= seq) return;
    if (*reinterpret_cast<volatile int*>(err)) return;
    const unsigned long long limit = *reinterpret_cast<const volatile unsigned long long*>(err + 2);
    const unsigned long long t0 = global_ns();
    while (ld_acquire_sys(flag) < seq) {
        if (global_ns() - t0 > limit) {
            *reinterpret_cast<volatile int*>(err) = 1;
            __threadfence();
            return;
        }
        __nanosleep(64);
    }
}

// ------------------------------ ghost cells ---------------------------------------
// Ghost value of one side from the adjacent (or periodic image) cell's 3 components.
// problem.py:761-768:  D: (target - 0.5*q + 0.0*0.0)/(0.5 - 0.0);  N: ((1-0.5)*q + 0.0*0.0)/(0.5-0.0)
__device__ __forceinline__ void ghost_val(const BcTable& bc, int side, const double v[3], double g[3]) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const int t = bc.type[side][c];
        const double a = v[bc.src[side][c]];
        if (t == GPF_BC_D)
            g[c] = (bc.val[side] - 0.5 * a + 0.0) / 0.5;
        else if (t == GPF_BC_N)
            g[c] = (0.5 * a + 0.0) / 0.5;
        else
            g[c] = v[c];
    }
}

// Calls f(gi, gj, g[3]) for every ghost cell whose final value derives from interior cell
// (i, j) holding v[3]; order of the reference: x sides first, then y sides over the full
// x range, so corners are y-fills of x-ghosts (problem.py:681-707).
template <class F>
__device__ __forceinline__ void for_each_ghost(const FvParams& P, int i, int j, const double v[3], F&& f) {
    const int Nx = P.L.Nx, Ny = P.L.Ny;
    if (!((i == 1) | (i == Nx) | (j == 1) | (j == Ny))) return;
    for (int s = 0; s < 2; ++s) {
        const int t0 = P.bc.type[s][0];
        if (t0 == GPF_BC_HALO) continue;
        const bool per = (t0 == GPF_BC_P);
        const int src = (s == 0) ? (per ? Nx : 1) : (per ? 1 : Nx);
        if (i != src) continue;
        double g[3];
        ghost_val(P.bc, s, v, g);
        const int gi = (s == 0) ? 0 : Nx + 1;
        f(gi, j, g);
        for (int t = 2; t < 4; ++t) {
            const bool pery = (P.bc.type[t][0] == GPF_BC_P);
            const int srcj = (t == 2) ? (pery ? Ny : 1) : (pery ? 1 : Ny);
            if (j != srcj) continue;
            double gg[3];
            ghost_val(P.bc, t, g, gg);
            f(gi, (t == 2) ? 0 : Ny + 1, gg);
        }
    }
    for (int t = 2; t < 4; ++t) {
        const bool pery = (P.bc.type[t][0] == GPF_BC_P);
        const int srcj = (t == 2) ? (pery ? Ny : 1) : (pery ? 1 : Ny);
        if (j != srcj) continue;
        double g[3];
        ghost_val(P.bc, t, v, g);
        f(i, (t == 2) ? 0 : Ny + 1, g);
    }
}

// ------------------------------ reductions -----------------------------------------
struct Red {
    double ekin;    // sum (jx^2 + jy^2)/rho/2          problem.py:342
    double vmax2;   // max (jx^2 + jy^2)/rho            problem.py:347 (sqrt is monotone)
    double cs2;     // max dp/drho                      stress.py:539
    int flags;      // 1: NaN in q, 2: rho < 0, 4 / 8: sqrt argument of v_max / v_sound NaN or negative
};
constexpr int FLAG_NAN = 1, FLAG_NEG = 2, FLAG_VNAN = 4, FLAG_CNAN = 8;
constexpr int FLAG_COMM = 16;   // slab exchange: a wait on a peer timed out somewhere (fv_kernels.cu: spin_until)

// kinetic energy and speed of one cell (problem.py:342, 347)
__device__ __forceinline__ void red_accum_v(Red& r, const double v[3]) {
    const double w = (v[1] * v[1] + v[2] * v[2]) / v[0];
    r.ekin += w / 2.0;
    if (!(w >= 0.0)) r.flags |= FLAG_VNAN;     // np.sqrt(negative) = NaN and np.max propagates it
    r.vmax2 = fmax(r.vmax2, w);
}
// squared sound speed of one cell (stress.py:539)
template <int EOS>
__device__ __forceinline__ void red_accum_c(Red& r, double rho, const EosParams& e) {
    const double c2 = eos_sound2<EOS>(rho, e);
    if (!(c2 >= 0.0)) r.flags |= FLAG_CNAN;
    r.cs2 = fmax(r.cs2, c2);
}
template <int EOS>
__device__ __forceinline__ void red_accum(Red& r, const double v[3], const EosParams& e) {
    red_accum_v(r, v);
    red_accum_c<EOS>(r, v[0], e);
}

// Boundary cells only (kept out of line so the interior path stays in registers): writes the
// ghost cells that derive from interior cell (i, j) and returns their share of the reductions.
//   qs[3]: value after this pass (before averaging), qn[3]: value stored for the cell.
// Corrector: validity is tested on the averaged array INCLUDING its averaged ghost cells
// (problem.py:563-565): ghost = (fill(q**) + ghost(q0)) / 2.
// X != nullptr (fused slab exchange, SlabX in fv_host.h): a cell of row 1 (row Nx) and its y-ghost images are also
// stored into the lower (upper) neighbour's ghost row of the same rotating buffer, over NVLink -- a neighbour's
// interior row with its y ghosts filled is exactly what the reference's x-then-y fill order leaves in an x-ghost row.
template <bool CORR, int EOS, bool SX = false>
__device__ __noinline__ Red boundary_cell(const FvParams& P, int i, int j, double s0, double s1, double s2,
                                          double n0, double n1, double n2, double* qout, const double* q0b,
                                          const SlabX* X = nullptr, int out_idx = 0) {
    Red r;
    r.ekin = 0.0; r.vmax2 = -INFINITY; r.cs2 = -INFINITY; r.flags = 0;
    const size_t plane = P.L.plane;
    const int pitch = P.L.pitch;
    if (CORR) {
        const double qs[3] = {s0, s1, s2};
        for_each_ghost(P, i, j, qs, [&](int gi, int gj, const double* gv) {
            const size_t a = (size_t)gi * pitch + (gj + 1);
#pragma unroll
            for (int cc = 0; cc < 3; ++cc) {
                const double av = (gv[cc] + q0b[cc * plane + a]) / 2.0;
                if (av != av) r.flags |= FLAG_NAN;
                if (cc == 0 && av < 0.0) r.flags |= FLAG_NEG;
            }
        });
    }
    const double qn[3] = {n0, n1, n2};
    auto remote = [&](int gj, const double* v) {          // (i, gj) of this slab -> the neighbours' ghost rows
#pragma unroll
        for (int sd = 0; sd < 2; ++sd) {
            double* pq = X->peer_q[sd][out_idx];
            if (pq && i == (sd == 0 ? 1 : P.L.Nx)) {
                const size_t a = (size_t)X->peer_row[sd] * pitch + (gj + 1);
                pq[a] = v[0];
                pq[X->peer_plane[sd] + a] = v[1];
                pq[2 * X->peer_plane[sd] + a] = v[2];
            }
        }
    };
    if constexpr (SX) remote(j, qn);
    for_each_ghost(P, i, j, qn, [&](int gi, int gj, const double* gv) {
        const size_t ga = (size_t)gi * pitch + (gj + 1);
        qout[ga] = gv[0];
        qout[plane + ga] = gv[1];
        qout[2 * plane + ga] = gv[2];
        if (CORR) red_accum<EOS>(r, gv, P.eos);
        if constexpr (SX) { if (gi == i) remote(gj, gv); }   // y-ghost image of an edge-row cell
    });
    return r;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

}  // namespace gpf
