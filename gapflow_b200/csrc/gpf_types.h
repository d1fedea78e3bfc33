// Internal types shared by the kernels and the C ABI (not part of the public header).
#pragma once
#include <stdint.h>
#include <stddef.h>
#include "../../include/gapflow_b200.h"

namespace gpf {

// EOS parameters, pre-combined on the host in the reference's own (Python float)
// operation order so device code only does the per-cell part.  See eos_pack() in api.cu.
struct EosParams {
    int kind;
    double p[10];
    double c[14];      // BWR temperature brackets
};

struct ViscParams {
    double U, V, eta, zeta, v1, v2, two_eta, two_v2;
};

// viscosity modifiers (stress.py:305-326; gpf_fv_set_viscosity)
struct GradAxis {           // np.gradient along one axis of the padded array
    int n;                  // points
    int uniform;            // NumPy found constant spacing: interior (f[i+1] - f[i-1]) / two_d
    double two_d;           // 2. * dx            (uniform)
    double d_first, d_last; // spacing of the one-sided differences at the two edges
    const double* abc;      // non-uniform: a[n], b[n], c[n] of  a f[i-1] + b f[i] + c f[i+1]  (device)
};
struct ViscModel {
    int piezo, thin;        // GPF_PIEZO_*, GPF_THIN_*
    int use_dens;           // Bayada-Chupin EOS: the piezo law takes the density (stress.py:307)
    double pp[4];           // piezo parameters (Roelands: pp[3] = log(mu0 / mu_inf), host)
    double tp[5];           // thinning parameters (Carreau: tp[4] = (N - 1) / a, host)
    GradAxis gx, gy;
};

// boundary table (device copy of gpf_fv_cfg.bc_*)
struct BcTable {
    int type[4][3];
    int src[4][3];
    double val[4];
    int edge[2];        // HALO x side whose ghost row is the global periodic ghost row (counted in reductions)
};

// tile geometry of the FV pass kernels
constexpr int TX = 8;               // tile rows (x, slow axis)
constexpr int TY = 64;              // tile cols (y, contiguous axis)
constexpr int BX = TX + 1;          // staged rows  (one upwind/downwind halo row)
constexpr int BY = TY + 4;          // staged cols: 2 left + TY + 2 right.  TMA needs the box origin on a
                                    // 16-byte boundary (even column) and a 16-byte multiple inner extent,
                                    // so the halo column on either side comes with one pad column
constexpr int FV_THREADS = 256;     // 2 cells per thread
// persistent CTAs per SM: the predictor's ring + closure planes take 69 KB (x-driven film), so 3 CTAs
// of 80 registers fit; the corrector also stages q0 (93 KB): 2 CTAs of 128 registers
#ifndef GPF_CTAS_PRED
#define GPF_CTAS_PRED 3
#endif
#ifndef GPF_CTAS_CORR
#define GPF_CTAS_CORR 2
#endif
__host__ __device__ constexpr int fv_ctas_per_sm(bool corr) { return corr ? GPF_CTAS_CORR : GPF_CTAS_PRED; }
// depth of the TMA ring per pass
#ifndef GPF_NST_PRED
#define GPF_NST_PRED 2
#endif
#ifndef GPF_NST_CORR
#define GPF_NST_CORR 2
#endif
__host__ __device__ constexpr int fv_nstage(bool corr) { return corr ? GPF_NST_CORR : GPF_NST_PRED; }

// storage layout: plane-major, rows padded.  Cell (i, j) of the reference's padded array
// (i in [0, Nx+1], j in [0, Ny+1]) lives at  comp*plane + i*pitch + (j + 1):
// one pad column in front so interior rows start on a 16-byte boundary.
struct Layout {
    int Nx, Ny;
    int pitch;          // doubles per row (multiple of 16)
    int rows;           // allocated rows (>= Nx+2, >= BX)
    size_t plane;       // doubles per component plane = rows*pitch
};

// device-resident loop state (one per solver); the finalize kernel advances it
struct DevState {
    long long step;
    double simtime, dt;
    double residual;
    double resbuf[5];
    int rescount;
    double ekin, ekin_old, vmax, vsound, dt_crit, cfl;
    int has_nan, has_neg, stop, converged;
    int halt;           // run-loop condition false: later steps are no-ops
    int cur;            // index of the buffer holding the current q (0..2)
    int comm_err;       // slab exchange: some rank's wait on a peer timed out; the run is stopped (see spin_until)
};

// per-block partial record written by reducing kernels, consumed by fv_finish_kernel
struct Partial {
    double ekin, vmax2, cs2, mass;
};

// one rank's folded record, all-gathered between ranks (slab decomposition)
struct Rec5 {
    double ekin, vmax2, cs2, mass, flags;
};

// Fused slab exchange (slab_comm.cuh): what a pass kernel needs to store its edge rows straight into the
// neighbours' ghost rows, to wait for theirs, and -- corrector -- to fold and publish this rank's record.
// Side 0 = the lower neighbour (it receives my row 1 into its row Nx_peer + 1), side 1 = the upper one (my row Nx
// into its row 0).
struct SlabX {
    int on;                               // 0: plain pass
    double* peer_q[2][3];                 // [side][rotating buffer]: the neighbour's q buffers (peer-mapped); null = no neighbour
    unsigned long long peer_plane[2];     // the neighbour's plane stride (doubles)
    int peer_row[2];                      // the ghost row of the neighbour that my edge row fills
    unsigned long long* peer_flag[2];     // the neighbour's flag that I raise when my edge row is complete
    const unsigned long long* my_flag;    // [2]: raised by my lower / upper neighbour
    unsigned long long* xseq;             // exchanges this rank has published (device-resident: graph replay)
    unsigned int* xdone;                  // CTA exit counter of a pass that the general kernel runs in full
    int* err;                             // communicator error block (spin_until)
    int edge_dup[2];                      // periodic ring: my row 1 (Nx) is also the global ghost row Nx_g + 1 (0) of the
                                          // reference's padded array; its cells count twice in the kinetic-energy sum
    // record fold (corrector): fv_record_send_kernel's job, done by the general launch
    unsigned char* const* peers;          // device table of every rank's communicator buffer
    unsigned long long off_recs, off_rflag;
    int world, rank;
    const unsigned long long* rseq;       // folds completed (advanced by fv_finish_kernel)
    int nrec;                             // per-CTA records of this rank's corrector
};

struct FvParams {
    Layout L;
    double dx, dy, rdx, rdy;
    ViscParams visc;
    EosParams eos;
    BcTable bc;
    double CFL, tol, dt0;
    int adaptive, max_it, mc_order;
    int has_extra;
    int visc_field;               // shear viscosity is a per-cell field (piezo / thinning), staged like the slip length
    int xflow;                    // V == 0 and dh/dy == 0 everywhere (x-driven film)
    int scalars_fast;             // dx, dy, viscosities, wall speeds, EOS parameters all in [2^-100, 2^100] or 0
    int static_fast;              // scalars_fast and every h in range, dh/dx, dh/dy, Ls in range or 0, Ls >= 0
    int nblk_x, nblk_y;           // tile grid (nblk_x tiles along y, nblk_y along x)
    int ncta;                     // persistent CTAs of a pass launch = per-CTA reduction records
};

}  // namespace gpf
