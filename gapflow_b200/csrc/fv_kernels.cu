// FV kernels + launchers.  Compile with -fmad=false (see closures.cuh).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "fv_pass.cuh"
#include "fv_host.h"
#include "slab_comm.cuh"
#include "closures_slip_gen.cuh"

namespace gpf {

// =====================================================================================
// Scalar bookkeeping kernels
// =====================================================================================
enum { FIN_STEP = 0, FIN_PRERUN = 1, FIN_PRERUN_KEEP_EKIN = 2, FIN_REFRESH = 3 };

// Reduces the per-block records in a fixed order (deterministic), then thread 0 advances
// the loop state the way Problem._post_update / _finalize / _pre_run do
// (problem.py:571-610, 428-443).
// Slab decomposition: `recs` (one Rec5 per rank, gathered in rank order) replaces the per-block
// records, so every rank folds the same numbers in the same order and advances identically.
// `rec_out` != nullptr: only fold this rank's per-block records into one Rec5 and return.
// ---- system-scope flag protocol of the peer-memory exchange (slab_comm.cuh) ----------------
__global__ void __launch_bounds__(256) fv_finish_kernel(const __grid_constant__ FvParams P, DevState* st,
                                                        const Partial* part, const int* pflags, int nblk,
                                                        int mode, int until_converged, RefreshOut* ro,
                                                        const Rec5* recs, int nrec, Rec5* rec_out, int xslot,
                                                        const unsigned long long* wait_flags = nullptr,
                                                        unsigned long long* seqp = nullptr, int* werr = nullptr,
                                                        int rec_stride = 5, double* recs_copy = nullptr) {
    const int tid = threadIdx.x;
    if (wait_flags) {        // records written into this rank's buffer by its peers (slab_comm.cuh)
        const unsigned long long seq = *seqp + 1;
        if (tid < nrec) spin_until(wait_flags + tid, seq, werr);
        __syncthreads();
        if (tid == 0) *seqp = seq;                       // this fold is consumed
        recs = reinterpret_cast<const Rec5*>(reinterpret_cast<const double*>(recs) + (size_t)(seq & 1ull) * nrec * rec_stride);
        if (recs_copy && tid < nrec)
            for (int k = 0; k < 5; ++k) recs_copy[tid * 5 + k] = __ldcg(reinterpret_cast<const double*>(recs) + (size_t)tid * rec_stride + k);
    }
    if (mode == FIN_STEP && (st->stop || (until_converged && st->halt))) return;
    const int cerr = (werr && *reinterpret_cast<volatile int*>(werr)) ? FLAG_COMM : 0;   // a wait of this rank timed out
    __shared__ double sE[256], sV[256], sC[256], sM[256];
    __shared__ int sF[256];
    double E = 0.0, VM = -INFINITY, CM = -INFINITY, M = 0.0;
    int F = 0;
    if (recs) {
        if (tid == 0)
            for (int r = 0; r < nrec; ++r) {      // rank order, sequential: identical on every rank
                const double* rr = reinterpret_cast<const double*>(recs) + (size_t)r * rec_stride;   // L2 reads: peers wrote them
                E += __ldcg(rr + 0);
                M += __ldcg(rr + 3);
                VM = fmax(VM, __ldcg(rr + 1));
                CM = fmax(CM, __ldcg(rr + 2));
                F |= (int)__ldcg(rr + 4);
            }
    } else {
        for (int b = tid; b < nblk; b += 256) {
            E += part[b].ekin;
            M += part[b].mass;
            VM = fmax(VM, part[b].vmax2);
            CM = fmax(CM, part[b].cs2);
            F |= pflags[b];
        }
        if (xslot >= 0 && tid == 255) {           // the ring-edge ghost rows' record (fv_edge_rows_kernel)
            E += part[xslot].ekin;
            M += part[xslot].mass;
            VM = fmax(VM, part[xslot].vmax2);
            CM = fmax(CM, part[xslot].cs2);
            F |= pflags[xslot];
        }
    }
    sE[tid] = E; sV[tid] = VM; sC[tid] = CM; sM[tid] = M; sF[tid] = F;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (tid < s) {
            sE[tid] += sE[tid + s];
            sM[tid] += sM[tid + s];
            sV[tid] = fmax(sV[tid], sV[tid + s]);
            sC[tid] = fmax(sC[tid], sC[tid + s]);
            sF[tid] |= sF[tid + s];
        }
        __syncthreads();
    }
    if (tid != 0) return;
    E = sE[0]; VM = sV[0]; CM = sC[0]; M = sM[0]; F = sF[0] | cerr;
    if (rec_out) {
        rec_out->ekin = E; rec_out->vmax2 = VM; rec_out->cs2 = CM; rec_out->mass = M; rec_out->flags = (double)F;
        return;
    }
    if (F & FLAG_COMM) {
        // some rank's wait on a peer timed out: the exchanged rows / records of this step are not trustworthy.
        // Every rank sees the flag in the same fold and stops; gpf_fv_get_scalars reports it.
        st->comm_err = 1;
        st->stop = 1;
        st->halt = 1;
        return;
    }
    const double qnan = nan("");
    // np.max propagates NaN (problem.py:347, stress.py:539)
    const double vmax = (F & FLAG_VNAN) ? qnan : sqrt(VM);
    const double vsound = (F & FLAG_CNAN) ? qnan : sqrt(CM);
    const double dxy = (P.dy < P.dx) ? P.dy : P.dx;          // Python min(dx, dy)

    if (mode == FIN_REFRESH) {
        ro->ekin = E; ro->vmax = vmax; ro->vsound = vsound; ro->mass = M;
        ro->has_nan = (F & FLAG_NAN) ? 1 : 0;
        ro->has_neg = (F & FLAG_NEG) ? 1 : 0;
        return;
    }
    if (mode == FIN_PRERUN || mode == FIN_PRERUN_KEEP_EKIN) {
        st->ekin = E; st->vmax = vmax; st->vsound = vsound;
        if (mode == FIN_PRERUN) st->ekin_old = E;              // problem.py:670
        st->dt_crit = dxy / (vmax + vsound);
        st->step = 0; st->simtime = 0.0; st->residual = 1.0;    // problem.py:429-432
        st->resbuf[0] = 1.0; st->rescount = 1;
        st->dt = P.adaptive ? P.CFL * st->dt_crit : P.dt0;      // problem.py:434-437
        st->cfl = st->dt / st->dt_crit;
        st->has_nan = 0; st->has_neg = 0; st->stop = 0;
        st->converged = (1.0 < P.tol) ? 1 : 0;
        st->halt = (st->converged || st->step >= P.max_it) ? 1 : 0;
        return;
    }
    // ---- FIN_STEP ----
    st->has_nan = (F & FLAG_NAN) ? 1 : 0;
    st->has_neg = (F & FLAG_NEG) ? 1 : 0;
    if (F & (FLAG_NAN | FLAG_NEG)) {
        // _finalize (problem.py:588-610): q0 is still the current buffer -> nothing to restore
        st->stop = 1;
        st->halt = 1;
        return;
    }
    st->cur = (st->cur + 2) % 3;
    st->ekin = E; st->vmax = vmax; st->vsound = vsound;
    st->dt_crit = dxy / (vmax + vsound);
    st->cfl = st->dt / st->dt_crit;
    const double res = fabs(E - st->ekin_old) / st->ekin_old / st->cfl;   // problem.py:578
    st->residual = res;
    if (st->rescount < 5) {
        st->resbuf[st->rescount++] = res;
    } else {
        for (int k = 0; k < 4; ++k) st->resbuf[k] = st->resbuf[k + 1];
        st->resbuf[4] = res;
    }
    st->ekin_old = E;
    st->step += 1;
    st->simtime += st->dt;
    if (P.adaptive) {
        st->dt = P.CFL * st->dt_crit;                                     // problem.py:585-586
        st->cfl = st->dt / st->dt_crit;
    }
    int conv = 1;
    for (int k = 0; k < st->rescount; ++k) conv &= (st->resbuf[k] < P.tol) ? 1 : 0;   // problem.py:360-362
    st->converged = conv;
    st->halt = (conv || st->step >= P.max_it) ? 1 : 0;
}

// Scalars of the current q over the whole padded array (problem.py:320-347).
template <int EOS>
__global__ void __launch_bounds__(256) fv_reduce_kernel(const __grid_constant__ FvParams P, const DevState* st,
                                                        double* const* qbufs, const double* topo,
                                                        Partial* part, int* pflags) {
    const double* q = qbufs[st->cur];
    const int Nx2 = P.L.Nx + 2, Ny2 = P.L.Ny + 2;
    const size_t plane = P.L.plane;
    Red red; red.ekin = 0.0; red.vmax2 = -INFINITY; red.cs2 = -INFINITY; red.flags = 0;
    double mass = 0.0;
    const size_t total = (size_t)Nx2 * Ny2;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (size_t)gridDim.x * blockDim.x) {
        const int i = (int)(idx / Ny2), j = (int)(idx % Ny2);
        // slab-internal halo rows are the neighbour rank's interior cells: counted there
        if ((i == 0 && P.bc.type[0][0] == GPF_BC_HALO && !P.bc.edge[0]) ||
            (i == Nx2 - 1 && P.bc.type[1][0] == GPF_BC_HALO && !P.bc.edge[1])) continue;
        const size_t a = (size_t)i * P.L.pitch + (j + 1);
        double v[3] = {q[a], q[plane + a], q[2 * plane + a]};
        if (v[0] != v[0] || v[1] != v[1] || v[2] != v[2]) red.flags |= FLAG_NAN;
        if (v[0] < 0.0) red.flags |= FLAG_NEG;
        red_accum<EOS>(red, v, P.eos);
        mass += v[0] * topo[a] * P.dx * P.dy;                 // problem.py:337
    }
    __shared__ double sred[8 * 4];
    __shared__ int sfl[8];
    double e = warp_sum(red.ekin), vm = warp_max(red.vmax2), cm = warp_max(red.cs2), ms = warp_sum(mass);
    int fl = red.flags;
#pragma unroll
    for (int oo = 16; oo > 0; oo >>= 1) fl |= __shfl_xor_sync(0xffffffffu, fl, oo);
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { sred[w * 4] = e; sred[w * 4 + 1] = vm; sred[w * 4 + 2] = cm; sred[w * 4 + 3] = ms; sfl[w] = fl; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double E = 0.0, VM = -INFINITY, CM = -INFINITY, M = 0.0; int F = 0;
        for (int ww = 0; ww < 8; ++ww) {
            E += sred[ww * 4]; VM = fmax(VM, sred[ww * 4 + 1]); CM = fmax(CM, sred[ww * 4 + 2]);
            M += sred[ww * 4 + 3]; F |= sfl[ww];
        }
        part[blockIdx.x].ekin = E; part[blockIdx.x].vmax2 = VM; part[blockIdx.x].cs2 = CM;
        part[blockIdx.x].mass = M; pflags[blockIdx.x] = F;
    }
}

// _communicate_ghost_buffers on the current q (problem.py:676-707): one thread per
// perimeter interior cell.
__global__ void fv_fill_ghosts_kernel(const __grid_constant__ FvParams P, const DevState* st, double* const* qbufs) {
    double* q = qbufs[st->cur];
    const int Nx = P.L.Nx, Ny = P.L.Ny;
    const int n = 2 * Nx + 2 * Ny;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    int i, j;
    if (t < Ny) { i = 1; j = t + 1; }
    else if (t < 2 * Ny) { i = Nx; j = t - Ny + 1; if (Nx == 1) return; }
    else if (t < 2 * Ny + Nx) { i = t - 2 * Ny + 1; j = 1; if (i == 1 || i == Nx) return; }
    else { i = t - 2 * Ny - Nx + 1; j = Ny; if (i == 1 || i == Nx || Ny == 1) return; }
    const size_t plane = P.L.plane;
    const size_t a = (size_t)i * P.L.pitch + (j + 1);
    const double v[3] = {q[a], q[plane + a], q[2 * plane + a]};
    for_each_ghost(P, i, j, v, [&](int gi, int gj, const double* gv) {
        const size_t ga = (size_t)gi * P.L.pitch + (gj + 1);
        q[ga] = gv[0]; q[plane + ga] = gv[1]; q[2 * plane + ga] = gv[2];
    });
}

// Slab decomposition along x (no counterpart in the reference, whose _communicate_ghost_buffers
// is a local copy, problem.py:676-707): rows 1 and Nx of buffer (cur + rel) % 3, y-ghost columns
// included, to/from a contiguous [side][comp][Ny+2] buffer.  A neighbour's interior row with its
// y ghosts filled is exactly what the reference's x-then-y fill order leaves in the x-ghost row.
__global__ void fv_halo_pack_kernel(const __grid_constant__ FvParams P, const DevState* st, double* const* qbufs,
                                    int rel, double* send) {
    const double* q = qbufs[(st->cur + rel) % 3];
    const int Ny2 = P.L.Ny + 2;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 6 * Ny2) return;
    const int side = t / (3 * Ny2), c = (t / Ny2) % 3, j = t % Ny2;
    const int i = side == 0 ? 1 : P.L.Nx;
    send[t] = q[c * P.L.plane + (size_t)i * P.L.pitch + (j + 1)];
}
__global__ void fv_halo_unpack_kernel(const __grid_constant__ FvParams P, const DevState* st, double* const* qbufs,
                                      int rel, const double* recv) {
    double* q = qbufs[(st->cur + rel) % 3];
    const int Ny2 = P.L.Ny + 2;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 6 * Ny2) return;
    const int side = t / (3 * Ny2), c = (t / Ny2) % 3, j = t % Ny2;
    if (P.bc.type[side][0] != GPF_BC_HALO) return;
    const int i = side == 0 ? 0 : P.L.Nx + 1;
    q[c * P.L.plane + (size_t)i * P.L.pitch + (j + 1)] = recv[t];
}

// The same exchange over peer memory (slab_comm.cuh): the sender stores its edge rows straight into
// the neighbours' buffers through NVLink and publishes a sequence number; the receiver waits for
// that number and fills its ghost rows.  One block each: they must fit beside the persistent CTAs
// of the interior launch that runs meanwhile.
// `done` counts finished blocks; the last one publishes (send) / advances the counter (receive).
__global__ void __launch_bounds__(256) fv_halo_send_kernel(const __grid_constant__ FvParams P, const DevState* st,
                                                           double* const* qbufs, int rel, double* dst_lo, double* dst_hi,
                                                           size_t par_stride, unsigned long long* flag_lo,
                                                           unsigned long long* flag_hi, const unsigned long long* seqp,
                                                           unsigned int* done) {
    const unsigned long long seq = *seqp + 1;             // this exchange; the receive kernel advances *seqp
    const double* q = qbufs[(st->cur + rel) % 3];
    const int Ny2 = P.L.Ny + 2;
    if (dst_lo) dst_lo += (seq & 1ull) * par_stride;
    if (dst_hi) dst_hi += (seq & 1ull) * par_stride;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < 6 * Ny2; t += gridDim.x * blockDim.x) {
        const int side = t / (3 * Ny2), rem = t % (3 * Ny2), c = rem / Ny2, j = rem % Ny2;
        double* dst = side == 0 ? dst_lo : dst_hi;       // row 1 -> lower neighbour, row Nx -> upper neighbour
        if (!dst) continue;
        const int i = side == 0 ? 1 : P.L.Nx;
        dst[rem] = q[c * P.L.plane + (size_t)i * P.L.pitch + (j + 1)];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();                           // cumulative: orders the whole block's stores
        if (atomicAdd(done, 1u) == gridDim.x - 1) {       // every block's stores are fenced
            *done = 0;
            __threadfence_system();
            if (flag_lo) st_release_sys(flag_lo, seq);
            if (flag_hi) st_release_sys(flag_hi, seq);
        }
    }
}
__global__ void __launch_bounds__(256) fv_halo_recv_kernel(const __grid_constant__ FvParams P, const DevState* st,
                                                           double* const* qbufs, int rel, const double* src, size_t par_stride,
                                                           const unsigned long long* flags, unsigned long long* seqp, int* err,
                                                           unsigned int* done) {
    const unsigned long long seq = *seqp + 1;
    if (threadIdx.x < 2 && P.bc.type[threadIdx.x][0] == GPF_BC_HALO) spin_until(flags + threadIdx.x, seq, err);
    __syncthreads();
    src += (seq & 1ull) * par_stride;
    double* q = qbufs[(st->cur + rel) % 3];
    const int Ny2 = P.L.Ny + 2;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < 6 * Ny2; t += gridDim.x * blockDim.x) {
        const int side = t / (3 * Ny2), c = (t / Ny2) % 3, j = t % Ny2;
        if (P.bc.type[side][0] != GPF_BC_HALO) continue;
        const int i = side == 0 ? 0 : P.L.Nx + 1;
        q[c * P.L.plane + (size_t)i * P.L.pitch + (j + 1)] = __ldcg(src + t);     // L2: a peer wrote it
    }
    __syncthreads();
    if (threadIdx.x == 0 && atomicAdd(done, 1u) == gridDim.x - 1) {              // every block has read *seqp
        *done = 0;
        *seqp = seq;
    }
}
// this rank's per-block records folded into one Rec5 (as fv_finish_kernel does for rec_out) and
// stored into every rank's record table, each followed by its sequence number
__global__ void __launch_bounds__(256) fv_record_send_kernel(const Partial* part, const int* pflags, int nblk, int xslot,
                                                             unsigned char* const* peers, size_t off_recs, size_t off_rflag,
                                                             int world, int rank, const unsigned long long* seqp, const int* err) {
    const unsigned long long seq = *seqp + 1;             // this fold; fv_finish_kernel advances *seqp
    const int parity = (int)(seq & 1ull);
    __shared__ double sE[256], sV[256], sC[256], sM[256];
    __shared__ int sF[256];
    const int tid = threadIdx.x;
    double E = 0.0, VM = -INFINITY, CM = -INFINITY, M = 0.0;
    int F = 0;
    for (int b = tid; b < nblk; b += 256) {
        E += part[b].ekin;
        M += part[b].mass;
        VM = fmax(VM, part[b].vmax2);
        CM = fmax(CM, part[b].cs2);
        F |= pflags[b];
    }
    if (xslot >= 0 && tid == 255) {
        E += part[xslot].ekin;
        M += part[xslot].mass;
        VM = fmax(VM, part[xslot].vmax2);
        CM = fmax(CM, part[xslot].cs2);
        F |= pflags[xslot];
    }
    sE[tid] = E; sV[tid] = VM; sC[tid] = CM; sM[tid] = M; sF[tid] = F;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (tid < s) {
            sE[tid] += sE[tid + s];
            sM[tid] += sM[tid + s];
            sV[tid] = fmax(sV[tid], sV[tid + s]);
            sC[tid] = fmax(sC[tid], sC[tid + s]);
            sF[tid] |= sF[tid + s];
        }
        __syncthreads();
    }
    if (tid < world) {
        double* dst = reinterpret_cast<double*>(peers[tid] + off_recs) + ((size_t)parity * world + rank) * 8;
        const int cf = *reinterpret_cast<const volatile int*>(err) ? FLAG_COMM : 0;   // tell every rank
        dst[0] = sE[0]; dst[1] = sV[0]; dst[2] = sC[0]; dst[3] = sM[0]; dst[4] = (double)(sF[0] | cf);
        __threadfence_system();
        st_release_sys(reinterpret_cast<unsigned long long*>(peers[tid] + off_rflag) + rank, seq);
    }
}

// Periodic-x ring: the global ghost rows (row 0 on the first rank, row Nx+1 on the last) belong
// to the reference's padded array, so validity, kinetic energy and the maxima count them
// (problem.py:320-347).  They arrive by exchange after the corrector pass; this folds them into
// one extra per-block record (slot `slot`).  One block.
template <int EOS>
__global__ void __launch_bounds__(256) fv_edge_rows_kernel(const __grid_constant__ FvParams P, const DevState* st,
                                                           double* const* qbufs, int rel, Partial* part, int* pflags,
                                                           int slot) {
    const double* q = qbufs[(st->cur + rel) % 3];
    const int Ny2 = P.L.Ny + 2;
    const size_t plane = P.L.plane;
    Red red; red.ekin = 0.0; red.vmax2 = -INFINITY; red.cs2 = -INFINITY; red.flags = 0;
    for (int side = 0; side < 2; ++side) {
        if (!(P.bc.type[side][0] == GPF_BC_HALO && P.bc.edge[side])) continue;
        const int i = side == 0 ? 0 : P.L.Nx + 1;
        for (int j = threadIdx.x; j < Ny2; j += blockDim.x) {
            const size_t a = (size_t)i * P.L.pitch + (j + 1);
            double v[3] = {q[a], q[plane + a], q[2 * plane + a]};
            if (v[0] != v[0] || v[1] != v[1] || v[2] != v[2]) red.flags |= FLAG_NAN;
            if (v[0] < 0.0) red.flags |= FLAG_NEG;
            red_accum<EOS>(red, v, P.eos);
        }
    }
    __shared__ double sred[8 * 3];
    __shared__ int sfl[8];
    double e = warp_sum(red.ekin), vm = warp_max(red.vmax2), cm = warp_max(red.cs2);
    int fl = red.flags;
#pragma unroll
    for (int oo = 16; oo > 0; oo >>= 1) fl |= __shfl_xor_sync(0xffffffffu, fl, oo);
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { sred[w * 3] = e; sred[w * 3 + 1] = vm; sred[w * 3 + 2] = cm; sfl[w] = fl; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double E = 0.0, VM = -INFINITY, CM = -INFINITY; int F = 0;
        for (int ww = 0; ww < 8; ++ww) { E += sred[ww * 3]; VM = fmax(VM, sred[ww * 3 + 1]); CM = fmax(CM, sred[ww * 3 + 2]); F |= sfl[ww]; }
        part[slot].ekin = E; part[slot].vmax2 = VM; part[slot].cs2 = CM; part[slot].mass = 0.0; pflags[slot] = F;
    }
}

// ---- viscosity modifiers (row F11) -----------------------------------------------------
// piezoviscosity (viscosity.py:34-66, 144-259); `arg` = pressure, or density for the mixture laws
__device__ __forceinline__ double piezo_law(const ViscModel& M, double mu0, double arg) {
    switch (M.piezo) {
    case GPF_PIEZO_BARUS:        // :165  mu0 * exp(aB * p)
        return mu0 * exp(M.pp[0] * arg);
    case GPF_PIEZO_ROELANDS:     // :197  mu0 * exp(log(mu0 / mu_inf) * (-1 + (1 + p / p_ref)**z)); pp[3] = the log
        return mu0 * exp(M.pp[3] * (-1 + pow(1 + arg / M.pp[1], M.pp[2])));
    case GPF_PIEZO_DUKLER: {     // :227-228  pp = {eta_v, rho_l, rho_v}
        const double alpha = (arg - M.pp[1]) / (M.pp[2] - M.pp[1]);
        return alpha * M.pp[0] + (1 - alpha) * mu0;
    }
    case GPF_PIEZO_MCADAMS: {    // :257-259
        const double alpha = (arg - M.pp[1]) / (M.pp[2] - M.pp[1]);
        const double Mf = alpha * M.pp[2] / arg;
        return M.pp[0] * mu0 / (mu0 * Mf + M.pp[0] * (1 - Mf));
    }
    default:                     // :64  np.ones_like(p) * mu
        return 1.0 * mu0;
    }
}

// np.gradient along one axis at index i (f[i-1], f[i], f[i+1]; edge_order = 1)
__device__ __forceinline__ double grad_axis(const GradAxis& G, int i, double fm, double f0, double fp) {
    if (i == 0) return (fp - f0) / G.d_first;
    if (i == G.n - 1) return (f0 - fm) / G.d_last;
    if (G.uniform) return (fp - fm) / G.two_d;
    return G.abc[i] * fm + G.abc[G.n + i] * f0 + G.abc[2 * G.n + i] * fp;
}

// Shear viscosity of every cell of the padded array, as WallStress.update and BulkStress.update build it
// at the head of each pass (stress.py:305-326 == 429-450) from the pressure field of the pass's input q:
// piezo law, then mu0 * thinning factor at the average shear rate (viscosity.py:99-141; the reference
// passes geo['U'], geo['V'] as the two wall speeds u1, u2).  The pressure of the cell and of its four
// neighbours is recomputed from the density (never stored); one plane is written.
__global__ void __launch_bounds__(256) fv_viscosity_kernel(const __grid_constant__ FvParams P,
                                                           const __grid_constant__ ViscModel M, const DevState* st,
                                                           double* const* qbufs, const double* topo, double* eta,
                                                           int rel) {
    const double* q0 = qbufs[(st->cur + rel) % 3];
    const int Nx2 = P.L.Nx + 2, Ny2 = P.L.Ny + 2;
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)Nx2 * Ny2) return;
    const int i = (int)(idx / Ny2), j = (int)(idx % Ny2);
    const size_t a = (size_t)i * P.L.pitch + (j + 1);
    const double rho = q0[a];
    const double p = eos_pressure<-1>(rho, P.eos);
    const double shear = P.visc.eta;
    double mu0 = shear;
    if (M.piezo != GPF_PIEZO_ABSENT) mu0 = piezo_law(M, shear, M.use_dens ? rho : p);
    double out = mu0;
    if (M.thin != GPF_THIN_ABSENT) {
        const double pxm = (i > 0) ? eos_pressure<-1>(q0[a - P.L.pitch], P.eos) : 0.0;
        const double pxp = (i < Nx2 - 1) ? eos_pressure<-1>(q0[a + P.L.pitch], P.eos) : 0.0;
        const double pym = (j > 0) ? eos_pressure<-1>(q0[a - 1], P.eos) : 0.0;
        const double pyp = (j < Ny2 - 1) ? eos_pressure<-1>(q0[a + 1], P.eos) : 0.0;
        const double gp = hypot(grad_axis(M.gx, i, pxm, p, pxp), grad_axis(M.gy, j, pym, p, pyp));
        const double h = topo[a];
        const double du_pois = h * gp / (2 * mu0);           // viscosity.py:106
        const double du_carr = (P.visc.V - P.visc.U) / h;    // :107 with (u1, u2) = (U, V)  (stress.py:317-318)
        const double rate = (fabs(du_pois + du_carr) + fabs(-du_pois + du_carr)) / 2.;
        double fac = 1.0;                                    // viscosity.py:94
        if (M.thin == GPF_THIN_EYRING) {                     // :284-285  tp = {tauE}
            const double tau0 = mu0 * rate;
            fac = M.tp[0] / tau0 * asinh(tau0 / M.tp[0]);
        } else if (M.thin == GPF_THIN_CARREAU) {             // :316-318  tp = {mu_inf, lam, a, N, (N - 1) / a}
            const double mu = M.tp[0] + (mu0 - M.tp[0]) * pow(1 + pow(M.tp[1] * rate, M.tp[2]), M.tp[4]);
            fac = mu / mu0;
        }
        out = mu0 * fac;
    }
    eta[a] = out;
}

// Closure fields as the reference stores them for sol.nc (stress.py:346-362, 452-459,
// 619-622): dense (ncomp, Nx+2, Ny+2) outputs; NULL outputs are skipped.
template <bool LS, int EOS>
__global__ void __launch_bounds__(256) fv_closures_kernel(const __grid_constant__ FvParams P, const DevState* st,
                                                          double* const* qbufs, const double* topo,
                                                          const double* extra, double* pr, double* wxz,
                                                          double* wyz, double* bulk, int rel,
                                                          const double* eta) {
    const double* q = qbufs[(st->cur + rel) % 3];
    const int Nx2 = P.L.Nx + 2, Ny2 = P.L.Ny + 2;
    const size_t np = (size_t)Nx2 * Ny2, plane = P.L.plane;
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= np) return;
    const int i = (int)(idx / Ny2), j = (int)(idx % Ny2);
    const size_t a = (size_t)i * P.L.pitch + (j + 1);
    Cell c;
    c.q0 = q[a]; c.q1 = q[plane + a]; c.q2 = q[2 * plane + a];
    c.h = topo[a]; c.hx = topo[plane + a]; c.hy = topo[2 * plane + a];
    c.Ls = LS ? extra[a] : 0.0;
    Visc V = reinterpret_cast<const Visc&>(P.visc);
    if (eta) {               // viscosity field (stress.py:305-326); v1, v2 per cell as viscous.py:82-83
        V.eta = eta[a];
        V.v1 = V.zeta + 4. / 3. * V.eta;
        V.v2 = V.zeta - 2. / 3. * V.eta;
        V.two_eta = 2 * V.eta;
        V.two_v2 = 2 * V.v2;
    }
    if (pr) pr[idx] = eos_pressure<EOS>(c.q0, P.eos);
    if (bulk) {
        double t0, t1, t2;
        stress_avg<LS>(c, V, t0, t1, t2);
        bulk[idx] = t0; bulk[np + idx] = t1; bulk[2 * np + idx] = t2;
    }
    if (wxz || wyz) {
        double b3, b4, u[6];
        stress_bottom<LS>(c, V, b3, b4);
        stress_top<LS, true>(c, V, u);
        // lower: comps 0-2 and 5 halved (zero for slip-top bottom wall); own shear in full
        for (int f = 0; f < 2; ++f) {
            double* w = f == 0 ? wxz : wyz;
            if (!w) continue;
            const int own = f == 0 ? 4 : 3;
            for (int k = 0; k < 12; ++k) w[k * np + idx] = 0.0;
            w[own * np + idx] = own == 4 ? b4 : b3;
            w[6 * np + idx] = u[0] / 2.0; w[7 * np + idx] = u[1] / 2.0; w[8 * np + idx] = u[2] / 2.0;
            w[11 * np + idx] = u[5] / 2.0;
            w[(own + 6) * np + idx] = u[own];
        }
    }
}

// ---- pure functions (B2 cut of the drop-in boundary) ---------------------------------
template <bool SOUND>
__global__ void eos_kernel(EosParams e, const double* dens, double* out, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = SOUND ? sqrt(eos_sound2<-1>(dens[i], e)) : eos_pressure<-1>(dens[i], e);
}

template <bool LS>
__global__ void stress_kernel(ViscParams vp, int which, const double* q, const double* h, const double* Ls,
                              double* out, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Cell c;
    c.q0 = q[i]; c.q1 = q[n + i]; c.q2 = q[2 * n + i];
    c.h = h[i]; c.hx = h[n + i]; c.hy = h[2 * n + i];
    c.Ls = LS ? Ls[i] : 0.0;
    const Visc& V = reinterpret_cast<const Visc&>(vp);
    if (which == 0) {          // stress_bottom
        double b3, b4;
        stress_bottom<LS>(c, V, b3, b4);
        for (int k = 0; k < 6; ++k) out[k * n + i] = 0.0;
        out[3 * n + i] = b3; out[4 * n + i] = b4;
    } else if (which == 1) {   // stress_top
        double u[6];
        stress_top<LS, true>(c, V, u);
        for (int k = 0; k < 6; ++k) out[k * n + i] = u[k];
    } else {                   // stress_avg
        double t0, t1, t2;
        stress_avg<LS>(c, V, t0, t1, t2);
        out[i] = t0; out[n + i] = t1; out[2 * n + i] = t2;
    }
}

// general form of the same three functions: slip at the top wall or at both walls, in-plane gradients of
// (rho, jx, jy) given (viscous.py:37-786 with dqx, dqy, slip).  Closed forms derived in tools/derive_slip_closures.py.
__host__ __device__ inline void stress_ex_cell(int which, int both, const SlipIn& in, double* out) {
    if (both) { if (which == 0) slip_both_bottom(in, out); else if (which == 1) slip_both_top(in, out); else slip_both_avg(in, out); }
    else { if (which == 0) slip_top_bottom(in, out); else if (which == 1) slip_top_top(in, out); else slip_top_avg(in, out); }
}
__global__ void stress_ex_kernel(ViscParams vp, int which, int both, const double* q, const double* h, const double* Ls,
                                 const double* dqx, const double* dqy, double* out, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    SlipIn in;
    in.h = h[i]; in.hx = h[n + i]; in.hy = h[2 * n + i]; in.Ls = Ls ? Ls[i] : 0.0;
    in.U = vp.U; in.V = vp.V; in.eta = vp.eta; in.zeta = vp.zeta;
    in.q0 = q[i]; in.q1 = q[n + i]; in.q2 = q[2 * n + i];
    in.dx0 = dqx ? dqx[i] : 0.0; in.dx1 = dqx ? dqx[n + i] : 0.0; in.dx2 = dqx ? dqx[2 * n + i] : 0.0;
    in.dy0 = dqy ? dqy[i] : 0.0; in.dy1 = dqy ? dqy[n + i] : 0.0; in.dy2 = dqy ? dqy[2 * n + i] : 0.0;
    double o[6];
    stress_ex_cell(which, both, in, o);
    const int nc = which == 2 ? 3 : 6;
    for (int k = 0; k < nc; ++k) out[k * n + i] = o[k];
}

// =====================================================================================
// Host side
// =====================================================================================
std::atomic<long long> g_launches{0};

// GPF_DEBUG_SYNC=1: synchronise after every launch and name the kernel that faulted
static int debug_check(const char* name, cudaStream_t s) {
    static const bool on = getenv("GPF_DEBUG_SYNC") != nullptr;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess && on) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) {
        set_error(std::string("CUDA error in kernel ") + name + ": " + cudaGetErrorString(e));
        return GPF_E_CUDA;
    }
    return GPF_OK;
}

Layout make_layout(int Nx, int Ny) {
    Layout L;
    L.Nx = Nx; L.Ny = Ny;
    int pitch = ((Ny + 3 + 15) / 16) * 16;
    if (pitch < 80) pitch = 80;                 // >= staged box width
    L.pitch = pitch;
    L.rows = (Nx + 2 < BX + 1) ? BX + 1 : Nx + 2;
    L.plane = (size_t)L.rows * L.pitch;
    return L;
}

static const double BWR_X[32] = {
    0.8623085097507421, 2.976218765822098, -8.402230115796038, 0.1054136629203555,
    -0.8564583828174598, 1.582759470107601, 0.7639421948305453, 1.753173414312048,
    2.798291772190376e+03, -4.8394220260857657e-02, 0.9963265197721935, -3.698000291272493e+01,
    2.084012299434647e+01, 8.305402124717285e+01, -9.574799715203068e+02, -1.477746229234994e+02,
    6.398607852471505e+01, 1.603993673294834e+01, 6.805916615864377e+01, -2.791293578795945e+03,
    -6.245128304568454, -8.116836104958410e+03, 1.488735559561229e+01, -1.059346754655084e+04,
    -1.131607632802822e+02, -8.867771540418822e+03, -3.986982844450543e+01, -4.689270299917261e+03,
    2.593535277438717e+02, -2.694523589434903e+03, -7.218487631550215e+02, 1.721802063863269e+02};

// Combines EOS parameters exactly as the reference's Python float expressions do.
// (host compiler: no FMA contraction -- build.py passes -ffp-contract=off)
static int eos_pack(const gpf_fv_cfg& c, EosParams* e) {
    const double* a = c.eos_par;
    for (int k = 0; k < 10; ++k) e->p[k] = 0.0;
    for (int k = 0; k < 14; ++k) e->c[k] = 0.0;
    e->kind = c.eos;
    switch (c.eos) {
    case GPF_EOS_DH:      // rho0, P0, C1, C2
        e->p[0] = a[0]; e->p[1] = a[1]; e->p[2] = a[2]; e->p[3] = a[3];
        e->p[4] = 0.99 * a[3] * a[0];
        e->p[5] = a[2] * a[0] * (a[3] - 1.0);
        e->p[6] = a[3] * a[0];
        e->p[9] = 1.0 / a[0];                 // RN(1/rho0) for div_fast
        return GPF_OK;
    case GPF_EOS_PL:      // rho0, P0, alpha
        e->p[0] = a[0]; e->p[1] = a[1]; e->p[2] = a[2];
        e->p[3] = 1. / (1. - 0.5 * a[2]);
        e->p[5] = -2.0 * a[1];
        e->p[6] = -2.0 / (a[2] - 2.0);
        e->p[7] = a[2] - 2;
        return GPF_OK;
    case GPF_EOS_VDW: {   // M, T, a, b
        const double R = 8.31446261815324;
        e->p[0] = a[0]; e->p[1] = R * a[1]; e->p[2] = a[2] / 10.; e->p[3] = a[3] / 1000.;
        e->p[4] = 2. * e->p[2];
        return GPF_OK;
    }
    case GPF_EOS_MT:      // rho0, P0, K, n
        e->p[0] = a[0]; e->p[1] = a[1]; e->p[2] = a[2] / a[3]; e->p[3] = a[3];
        e->p[5] = a[2] / pow(a[0], a[3]);
        e->p[6] = a[3] - 1;
        return GPF_OK;
    case GPF_EOS_CUBIC:   // a, b, c, d
        e->p[0] = a[0]; e->p[1] = a[1]; e->p[2] = a[2]; e->p[3] = a[3];
        e->p[4] = 3 * a[0]; e->p[5] = 2 * a[1];
        return GPF_OK;
    case GPF_EOS_BWR: {   // T, gamma
        const double T = a[0], T2 = pow(T, 2), T3 = pow(T, 3), T4 = pow(T, 4);   // Python float **
        const double* x = BWR_X;
        e->p[0] = T; e->p[1] = a[1];
        e->c[0] = x[0] * T + x[1] * sqrt(T) + x[2] + x[3] / T + x[4] / T2;
        e->c[1] = x[5] * T + x[6] + x[7] / T + x[8] / T2;
        e->c[2] = x[9] * T + x[10] + x[11] / T;
        e->c[3] = x[12];
        e->c[4] = x[13] / T + x[14] / T2;
        e->c[5] = x[15] / T;
        e->c[6] = x[16] / T + x[17] / T2;
        e->c[7] = x[18] / T2;
        e->c[8] = x[19] / T2 + x[20] / T3;
        e->c[9] = x[21] / T2 + x[22] / T4;
        e->c[10] = x[23] / T2 + x[24] / T3;
        e->c[11] = x[25] / T2 + x[26] / T4;
        e->c[12] = x[27] / T2 + x[28] / T3;
        e->c[13] = x[29] / T2 + x[30] / T3 + x[31] / T4;
        return GPF_OK;
    }
    case GPF_EOS_BAYADA: {   // rho_l, rho_v, c_l, c_v
        const double rl = a[0], rv = a[1], cl2 = pow(a[2], 2), cv2 = pow(a[3], 2);
        const double rv2 = pow(rv, 2), rl2 = pow(rl, 2);
        const double N = rv * cv2 * rl * cl2 * (rv - rl) / (rv2 * cv2 - rl2 * cl2);
        const double Pcav = rv * cv2 - N * log(rv2 * cv2 / (rl2 * cl2));
        const double cc = a[3] * a[2];
        e->p[0] = rl; e->p[1] = rv; e->p[2] = cl2; e->p[3] = cv2; e->p[4] = N; e->p[5] = Pcav;
        e->p[6] = rv * cv2; e->p[7] = rl * cl2; e->p[8] = rv * rl * pow(cc, 2);
        return GPF_OK;
    }
    }
    set_error("unknown EOS id");
    return GPF_E_INVALID;
}

// persistent CTAs of a pass launch: fv_ctas_per_sm() per SM of the current device minus `reserve`,
// at most one per tile -- and then the largest count G below that whose column stride G mod nty is
// coprime with nty (nty = tiles along y).  CTA b walks tiles b, b + G, b + 2G, ...; tile t sits in
// tile column t mod nty.  When G mod nty shares a factor >= 4 with nty every CTA keeps returning to
// the same few tile columns, and on B200 that costs 15-20 % of the pass (4096^2, nty = 64: 296 CTAs
// 0.310 ms, 295 CTAs 0.256 ms per corrector pass; 444 / 443: 0.210 / 0.194 ms per predictor pass;
// every multiple of 4 is slow, everything else fast: profiles/r1r_grid_scan.txt).
static int gcd_int(int a, int b) { while (b) { const int t = a % b; a = b; b = t; } return a; }
static int pass_ctas(int ntiles, bool corr, int nty = 0, int reserve = 0) {
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0)
            sms = 148;            // B200 (also used when no device is present: workspace sizing only)
        cudaGetLastError();
    }
    int n = sms * fv_ctas_per_sm(corr);
    if (n > 4 * reserve) n -= reserve;
    if (ntiles <= n) return ntiles;              // one tile per CTA: no walk
    if (nty > 1)
        for (int g = n; g > n - 16 && g > 1; --g)
            if (gcd_int(g % nty, nty) == 1) return g;
    return n;
}

int pack_params(const gpf_fv_cfg& c, FvParams* P) {
    if (c.Nx < 1 || c.Ny < 1) { set_error("Nx, Ny must be >= 1"); return GPF_E_INVALID; }
    if (c.mc_order != 0 && c.mc_order != 1 && c.mc_order != -1) { set_error("MC_order must be -1, 0 or 1"); return GPF_E_INVALID; }
    if (c.arith != GPF_ARITH_EXACT && c.arith != GPF_ARITH_FAST) { set_error("arith must be GPF_ARITH_EXACT or GPF_ARITH_FAST"); return GPF_E_INVALID; }
    P->L = make_layout(c.Nx, c.Ny);
    P->dx = c.dx; P->dy = c.dy;
    P->rdx = 1.0 / c.dx; P->rdy = 1.0 / c.dy;     // RN reciprocals for div_fast
    ViscParams& v = P->visc;
    v.U = c.U; v.V = c.V; v.eta = c.eta; v.zeta = c.zeta;
    v.v1 = c.zeta + 4. / 3. * c.eta;          // Python: zeta + 4 / 3 * eta
    v.v2 = c.zeta - 2. / 3. * c.eta;
    v.two_eta = 2 * c.eta;
    v.two_v2 = 2 * v.v2;
    int rc = eos_pack(c, &P->eos);
    if (rc) return rc;
    for (int s = 0; s < 4; ++s) {
        P->bc.val[s] = c.bc_val[s];
        for (int k = 0; k < 3; ++k) {
            const int t = c.bc_type[s][k];
            if (t < 0 || t > 3 || c.bc_src[s][k] < 0 || c.bc_src[s][k] > 2) { set_error("bad boundary table"); return GPF_E_INVALID; }
            const int t0 = c.bc_type[s][0];
            const bool whole = (t0 == GPF_BC_P || t0 == GPF_BC_HALO);   // side-wide kinds
            if (whole ? (t != t0) : !(t == GPF_BC_D || t == GPF_BC_N)) {
                set_error("a side mixes periodic with D/N components: the reference leaves such ghost cells undefined");
                return GPF_E_INVALID;
            }
            if (s >= 2 && t == GPF_BC_HALO) { set_error("HALO sides are x-only (slab decomposition along x)"); return GPF_E_INVALID; }
            P->bc.type[s][k] = t; P->bc.src[s][k] = c.bc_src[s][k];
        }
    }
    for (int s = 0; s < 2; ++s) P->bc.edge[s] = (c.bc_type[s][0] == GPF_BC_HALO && c.halo_edge[s]) ? 1 : 0;
    {
        auto okr = [](double x) { return x == 0.0 || (fabs(x) >= 0x1p-100 && fabs(x) < 0x1p100); };
        bool ok = okr(c.dx) && c.dx != 0.0 && okr(c.dy) && c.dy != 0.0 && okr(v.U) && okr(v.V) && okr(v.eta) &&
                  okr(v.zeta) && okr(v.v1) && okr(v.v2);
        for (int k = 0; k < 10; ++k) ok = ok && okr(P->eos.p[k]);
        P->scalars_fast = ok ? 1 : 0;
    }
    P->CFL = c.CFL; P->tol = c.tol; P->dt0 = c.dt;
    P->adaptive = c.adaptive; P->max_it = c.max_it; P->mc_order = c.mc_order;
    P->has_extra = c.has_extra;
    P->visc_field = c.visc_field ? 1 : 0;
    P->xflow = 0;                 // decided when the topography is uploaded
    P->static_fast = 0;           // decided when topography / slip length are uploaded
    P->nblk_y = (c.Nx + TX - 1) / TX;
    P->nblk_x = (c.Ny + TY - 1) / TY;
    P->ncta = pass_ctas(P->nblk_x * P->nblk_y, true, P->nblk_x);      // the corrector writes the records
    return GPF_OK;
}

__global__ void any_nonzero_kernel(const double* plane, int pitch, int nx2, int ny2, int* flag) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)nx2 * ny2) return;
    const int i = (int)(idx / ny2), j = (int)(idx % ny2);
    if (plane[(size_t)i * pitch + (j + 1)] != 0.0) *flag = 1;
}

// flag |= 1 if any value of the plane lies outside the straight-line division range
// (nz: zero not allowed; nonneg: negative not allowed)
__global__ void range_check_kernel(const double* plane, int pitch, int nx2, int ny2, int nz, int nonneg, int* flag) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)nx2 * ny2) return;
    const int i = (int)(idx / ny2), j = (int)(idx % ny2);
    const double v = plane[(size_t)i * pitch + (j + 1)];
    const bool ok = (nz ? in_fast_range_nz(v) : in_fast_range(v)) && !(nonneg && __double2hiint(v) < 0 && v != 0.0);
    if (!ok) *flag = 1;
}

// ---- TMA descriptors ----------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// planes of `ncomp` x rows x pitch doubles; box = (by, bx, nbox) (default (BY, BX, ncomp))
static int make_map(CUtensorMap* m, double* base, const Layout& L, int ncomp, int by = BY, int bx = BX, int nbox = 0) {
    EncodeTiledFn enc = get_encode();
    if (!enc) { set_error("cuTensorMapEncodeTiled not available from the driver"); return GPF_E_CUDA; }
    cuuint64_t dims[3] = {(cuuint64_t)L.pitch, (cuuint64_t)L.rows, (cuuint64_t)ncomp};
    cuuint64_t strides[2] = {(cuuint64_t)L.pitch * 8, (cuuint64_t)L.plane * 8};
    cuuint32_t box[3] = {(cuuint32_t)by, (cuuint32_t)bx, (cuuint32_t)(nbox ? nbox : ncomp)};
    cuuint32_t es[3] = {1, 1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, base, dims, strides, box, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed, CUresult " + std::to_string((int)r)); return GPF_E_CUDA; }
    return GPF_OK;
}

// ---- solver --------------------------------------------------------------------------
static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
static const int REDUCE_BLOCKS = 148 * 8;

size_t FvSolver::workspace_bytes(const gpf_fv_cfg& c) {
    Layout L = make_layout(c.Nx, c.Ny);
    size_t plane_b = align_up(L.plane * 8, 256);
    int nb = REDUCE_BLOCKS + 1;                        // >= persistent CTAs of a pass + the edge-row record
    size_t b = 0;
    b += 3 * 3 * plane_b;      // q x3 buffers
    b += 3 * plane_b;          // topo
    b += plane_b;              // extra
    if (c.visc_field)          // viscosity plane + np.gradient tables
        b += plane_b + align_up((size_t)3 * (c.Nx + 2) * 8, 256) + align_up((size_t)3 * (c.Ny + 2) * 8, 256);
    b += align_up(sizeof(DevState), 256) + align_up(sizeof(RefreshOut), 256);
    b += align_up((size_t)nb * sizeof(Partial), 256) + align_up((size_t)nb * sizeof(int), 256);
    b += 256 + 256;            // qbuf pointer table + slack for base alignment
    return b;
}

int FvSolver::init(const gpf_fv_cfg& c, void* ws, size_t bytes) {
    cfg = c;
    int rc = pack_params(c, &P);
    if (rc) return rc;
    if (bytes < workspace_bytes(c)) { set_error("workspace too small"); return GPF_E_INVALID; }
    GPF_CUDA(cudaGetDevice(&device));
    cudaPointerAttributes at;
    GPF_CUDA(cudaPointerGetAttributes(&at, ws));
    if (at.type != cudaMemoryTypeDevice && at.type != cudaMemoryTypeManaged) { set_error("workspace is not device memory"); return GPF_E_INVALID; }
    const Layout& L = P.L;
    size_t plane_b = align_up(L.plane * 8, 256);
    unsigned char* p = reinterpret_cast<unsigned char*>(align_up(reinterpret_cast<size_t>(ws), 256));
    GPF_CUDA(cudaMemset(ws, 0, bytes));
    for (int k = 0; k < 3; ++k) { qbuf[k] = reinterpret_cast<double*>(p); p += 3 * plane_b; }
    topo = reinterpret_cast<double*>(p); p += 3 * plane_b;
    extra = reinterpret_cast<double*>(p); p += plane_b;
    if (c.visc_field) {
        eta = reinterpret_cast<double*>(p); p += plane_b;
        gradtab = reinterpret_cast<double*>(p);
        p += align_up((size_t)3 * (c.Nx + 2) * 8, 256) + align_up((size_t)3 * (c.Ny + 2) * 8, 256);
    }
    visc_set = false;
    st = reinterpret_cast<DevState*>(p); p += align_up(sizeof(DevState), 256);
    refresh = reinterpret_cast<RefreshOut*>(p); p += align_up(sizeof(RefreshOut), 256);
    npartial = REDUCE_BLOCKS + 1;
    if (P.ncta + 1 > npartial) { set_error("device has more SMs than the record table allows"); return GPF_E_INVALID; }
    partial = reinterpret_cast<Partial*>(p); p += align_up((size_t)npartial * sizeof(Partial), 256);
    pflags = reinterpret_cast<int*>(p); p += align_up((size_t)npartial * sizeof(int), 256);
    qtable = reinterpret_cast<double**>(p);
    redo = reinterpret_cast<int*>(p + 128);         // zeroed with the workspace
    p += 256;
    // planes are allocated with stride plane_b; the layout's plane stride must match
    P.L.plane = plane_b / 8;
    GPF_CUDA(cudaMemcpy(qtable, qbuf, sizeof(qbuf), cudaMemcpyHostToDevice));
    for (int k = 0; k < 3; ++k) if ((rc = make_map(&qmap[k], qbuf[k], P.L, 3))) return rc;
    for (int k = 0; k < 3; ++k) if ((rc = make_map(&q0map[k], qbuf[k], P.L, 3, TY, TX))) return rc;
    if ((rc = make_map(&topo_map, topo, P.L, 3))) return rc;
    if ((rc = make_map(&topo2_map, topo, P.L, 3, BY, BX, 2))) return rc;
    if ((rc = make_map(&ls_map, extra, P.L, 1))) return rc;
    if (eta && (rc = make_map(&eta_map, eta, P.L, 1))) return rc;
    topo_fast = 0;
    ls_fast = 1;               // the slip-length plane starts as zeros
    // _initialize (problem.py:662-670): rho0, rho0*U/2, rho0*V/2 everywhere incl. ghosts
    {
        const size_t np = (size_t)(c.Nx + 2) * (c.Ny + 2);
        std::vector<double> h(3 * np);
        const double v[3] = {c.rho0, c.rho0 * c.U / 2.0, c.rho0 * c.V / 2.0};
        for (int k = 0; k < 3; ++k) for (size_t t = 0; t < np; ++t) h[k * np + t] = v[k];
        rc = upload(GPF_F_Q, h.data(), 3, 0, 0);
        if (rc) return rc;
        GPF_CUDA(cudaStreamSynchronize(0));
    }
    DevState s0;
    memset(&s0, 0, sizeof(s0));
    s0.dt = c.dt; s0.residual = 1.0; s0.resbuf[0] = 1.0; s0.rescount = 1;
    GPF_CUDA(cudaMemcpy(st, &s0, sizeof(s0), cudaMemcpyHostToDevice));
    dirty_scalars = true;
    refresh_valid = false;
    timing = false;
    ev_steps = 0;
    return GPF_OK;
}

void FvSolver::drop_graphs() {
    for (auto& g : step_graph)
        if (g) { cudaGraphExecDestroy(g); g = nullptr; }
    graph_stale = true;          // a slab-step graph held by a communicator is rebuilt on its next use
}

int FvSolver::destroy() {
    for (auto e : ev) cudaEventDestroy(e);
    ev.clear();
    for (auto e : cp_event) cudaEventDestroy(e);
    cp_event.clear();
    for (int k = 0; k < 2; ++k)
        if (cp_stream[k]) { cudaStreamDestroy(cp_stream[k]); cp_stream[k] = nullptr; }
    drop_graphs();
    return GPF_OK;
}

double* FvSolver::field_ptr(int field, int* nc) {
    switch (field) {
    case GPF_F_TOPO: *nc = 3; return topo;
    case GPF_F_EXTRA: *nc = 1; return extra;
    }
    *nc = 0;
    return nullptr;
}

int FvSolver::read_cur(cudaStream_t s, int* cur) {
    DevState h;
    GPF_CUDA(cudaMemcpyAsync(&h, st, sizeof(h), cudaMemcpyDeviceToHost, s));
    GPF_CUDA(cudaStreamSynchronize(s));
    *cur = h.cur;
    return GPF_OK;
}

int FvSolver::upload(int field, const double* data, int ncomp, int on_device, cudaStream_t s) {
    int ncmax = 3;
    double* base;
    if (field == GPF_F_Q) {
        int cur; int rc = read_cur(s, &cur); if (rc) return rc;
        base = qbuf[cur];
        if (ncomp != 3) { set_error("q has 3 components"); return GPF_E_INVALID; }
    } else {
        base = field_ptr(field, &ncmax);
        if (!base) { set_error("unknown field id"); return GPF_E_INVALID; }
        if (ncomp < ncmax || ncomp > ncmax + 1) { set_error("wrong component count"); return GPF_E_INVALID; }
    }
    const size_t Ny2 = cfg.Ny + 2, Nx2 = cfg.Nx + 2;
    for (int k = 0; k < ncmax; ++k)
        GPF_CUDA(cudaMemcpy2DAsync(base + k * P.L.plane + 1, (size_t)P.L.pitch * 8, data + (size_t)k * Nx2 * Ny2,
                                   Ny2 * 8, Ny2 * 8, Nx2, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, s));
    if (field == GPF_F_Q) { dirty_scalars = true; refresh_valid = false; }
    if (field == GPF_F_TOPO || field == GPF_F_EXTRA) {
        // static inputs are classified once: x-driven film specialisation (V == 0 and dh/dy == 0
        // everywhere: exact zeros drop out) and the straight-line division range
        int* flag = reinterpret_cast<int*>(refresh);
        GPF_CUDA(cudaMemsetAsync(flag, 0, 2 * sizeof(int), s));
        const size_t np = Nx2 * Ny2;
        const int blocks = (int)((np + 255) / 256);
        if (field == GPF_F_TOPO) {
            any_nonzero_kernel<<<blocks, 256, 0, s>>>(topo + 2 * P.L.plane, P.L.pitch, (int)Nx2, (int)Ny2, flag);
            range_check_kernel<<<blocks, 256, 0, s>>>(topo, P.L.pitch, (int)Nx2, (int)Ny2, 1, 0, flag + 1);
            range_check_kernel<<<blocks, 256, 0, s>>>(topo + P.L.plane, P.L.pitch, (int)Nx2, (int)Ny2, 0, 0, flag + 1);
            range_check_kernel<<<blocks, 256, 0, s>>>(topo + 2 * P.L.plane, P.L.pitch, (int)Nx2, (int)Ny2, 0, 0, flag + 1);
            g_launches += 4;
        } else {
            range_check_kernel<<<blocks, 256, 0, s>>>(extra, P.L.pitch, (int)Nx2, (int)Ny2, 0, 1, flag + 1);
            g_launches++;
        }
        GPF_CUDA(cudaGetLastError());
        int h[2] = {1, 1};
        GPF_CUDA(cudaMemcpyAsync(h, flag, 2 * sizeof(int), cudaMemcpyDeviceToHost, s));
        GPF_CUDA(cudaStreamSynchronize(s));
        if (field == GPF_F_TOPO) {
            P.xflow = (h[0] == 0 && cfg.V == 0.0) ? 1 : 0;
            topo_fast = h[1] == 0;
        } else {
            ls_fast = h[1] == 0;
        }
        P.static_fast = (P.scalars_fast && topo_fast && ls_fast) ? 1 : 0;
        drop_graphs();            // the captured launches carry P by value
    }
    return GPF_OK;
}

int FvSolver::download(int field, double* data, int ncomp, int on_device, cudaStream_t s) {
    int ncmax = 3;
    const double* base;
    if (field == GPF_F_Q) {
        int cur; int rc = read_cur(s, &cur); if (rc) return rc;
        base = qbuf[cur];
    } else {
        base = field_ptr(field, &ncmax);
        if (!base) { set_error("unknown field id"); return GPF_E_INVALID; }
    }
    if (ncomp < ncmax) { set_error("wrong component count"); return GPF_E_INVALID; }
    const size_t Ny2 = cfg.Ny + 2, Nx2 = cfg.Nx + 2;
    for (int k = 0; k < ncmax; ++k)
        GPF_CUDA(cudaMemcpy2DAsync(data + (size_t)k * Nx2 * Ny2, Ny2 * 8, base + k * P.L.plane + 1, (size_t)P.L.pitch * 8,
                                   Ny2 * 8, Nx2, on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, s));
    return GPF_OK;
}

// the pass kernels live in fv_pass_exact.cu / fv_pass_fast.cu (same source, two arithmetic contracts)
int launch_pass_exact(const PassArgs& A, bool corr, bool ls, bool dh, bool xf, bool vf, int n, cudaStream_t s);
int launch_pass_fast(const PassArgs& A, bool corr, bool ls, bool dh, bool xf, bool vf, int n, cudaStream_t s);

// ---- viscosity modifiers: host side ----------------------------------------------------
// np.gradient's per-axis set-up (numpy/lib/_function_base_impl.py, gradient(): `distances` given as
// coordinates): diffx = np.diff(x); constant spacing (diffx == diffx[0]).all() collapses to the scalar
// form, else the three coefficient arrays of the second-order non-uniform stencil.  Host arithmetic is
// compiled with -ffp-contract=off: same roundings as NumPy's.
static void grad_axis_setup(const double* x, int n, GradAxis* G, std::vector<double>* abc) {
    std::vector<double> d(n - 1);
    for (int i = 0; i + 1 < n; ++i) d[i] = x[i + 1] - x[i];
    bool uni = true;
    for (int i = 0; i + 1 < n; ++i) uni = uni && (d[i] == d[0]);
    G->n = n;
    G->uniform = uni ? 1 : 0;
    G->two_d = 2. * d[0];
    G->d_first = d[0];
    G->d_last = uni ? d[0] : d[n - 2];
    abc->assign((size_t)3 * n, 0.0);
    if (!uni)
        for (int i = 1; i + 1 < n; ++i) {
            const double dx1 = d[i - 1], dx2 = d[i];
            (*abc)[i] = -(dx2) / (dx1 * (dx1 + dx2));
            (*abc)[n + i] = (dx2 - dx1) / (dx1 * dx2);
            (*abc)[2 * n + i] = dx1 / (dx2 * (dx1 + dx2));
        }
}

int FvSolver::set_viscosity(int piezo, const double* pp, int thin, const double* tp, const double* x, const double* y) {
    if (!P.visc_field) { set_error("gpf_fv_set_viscosity: the solver was created with cfg.visc_field = 0"); return GPF_E_STATE; }
    if (piezo < GPF_PIEZO_ABSENT || piezo > GPF_PIEZO_MCADAMS || thin < GPF_THIN_ABSENT || thin > GPF_THIN_CARREAU) {
        set_error("unknown piezoviscosity / shear-thinning law id"); return GPF_E_INVALID;
    }
    if (piezo == GPF_PIEZO_ABSENT && thin == GPF_THIN_ABSENT) { set_error("visc_field set but neither block present"); return GPF_E_INVALID; }
    for (int sd = 0; sd < 2; ++sd)
        if (P.bc.type[sd][0] == GPF_BC_HALO) {
            set_error("viscosity modifiers on a slab (GPF_BC_HALO) solver: the pressure gradient at a halo row needs a "
                      "second halo row; not supported");
            return GPF_E_INVALID;
        }
    memset(&VM, 0, sizeof(VM));
    VM.piezo = piezo; VM.thin = thin;
    VM.use_dens = (P.eos.kind == GPF_EOS_BAYADA) ? 1 : 0;
    const int npp = (piezo == GPF_PIEZO_BARUS) ? 1 : (piezo >= GPF_PIEZO_ROELANDS ? 3 : 0);
    const int ntp = (thin == GPF_THIN_EYRING) ? 1 : (thin == GPF_THIN_CARREAU ? 4 : 0);
    if ((npp && !pp) || (ntp && !tp)) { set_error("law parameters missing"); return GPF_E_INVALID; }
    for (int k = 0; k < npp; ++k) VM.pp[k] = pp[k];
    for (int k = 0; k < ntp; ++k) VM.tp[k] = tp[k];
    if (piezo == GPF_PIEZO_ROELANDS) VM.pp[3] = log(cfg.eta / VM.pp[0]);
    if (thin == GPF_THIN_CARREAU) VM.tp[4] = (VM.tp[3] - 1) / VM.tp[2];
    const int nx = cfg.Nx + 2, ny = cfg.Ny + 2;
    if (thin != GPF_THIN_ABSENT) {
        if (!x || !y) { set_error("shear thinning needs the cell-centre coordinates x, y"); return GPF_E_INVALID; }
        std::vector<double> ax, ay;
        grad_axis_setup(x, nx, &VM.gx, &ax);
        grad_axis_setup(y, ny, &VM.gy, &ay);
        double* dx_tab = gradtab;
        double* dy_tab = gradtab + align_up((size_t)3 * nx * 8, 256) / 8;
        GPF_CUDA(cudaMemcpy(dx_tab, ax.data(), ax.size() * 8, cudaMemcpyHostToDevice));
        GPF_CUDA(cudaMemcpy(dy_tab, ay.data(), ay.size() * 8, cudaMemcpyHostToDevice));
        VM.gx.abc = dx_tab;
        VM.gy.abc = dy_tab;
    }
    visc_set = true;
    drop_graphs();
    return GPF_OK;
}

int FvSolver::launch_viscosity(int rel, cudaStream_t s) {
    if (!visc_set) { set_error("cfg.visc_field is set: call gpf_fv_set_viscosity() before stepping"); return GPF_E_STATE; }
    const size_t np = (size_t)(cfg.Nx + 2) * (cfg.Ny + 2);
    fv_viscosity_kernel<<<(int)((np + 255) / 256), 256, 0, s>>>(P, VM, st, qtable, topo, eta, rel);
    g_launches++;
    return debug_check("fv_viscosity_kernel", s);
}

int FvSolver::viscosity(int which, double* out, int on_device, cudaStream_t s) {
    if (!P.visc_field) { set_error("gpf_fv_viscosity: the solver has no viscosity field"); return GPF_E_STATE; }
    if (which < 0 || which > 1 || !out) { set_error("gpf_fv_viscosity: bad argument"); return GPF_E_INVALID; }
    int rc = launch_viscosity(which ? 2 : 0, s);      // after a step the predictor's output sits two buffers on
    if (rc) return rc;
    const size_t Ny2 = cfg.Ny + 2, Nx2 = cfg.Nx + 2;
    GPF_CUDA(cudaMemcpy2DAsync(out, Ny2 * 8, eta + 1, (size_t)P.L.pitch * 8, Ny2 * 8, Nx2,
                               on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, s));
    if (!on_device) GPF_CUDA(cudaStreamSynchronize(s));
    return GPF_OK;
}

// part: GPF_PART_ALL every tile row; GPF_PART_EDGE the first and the last tile row (whose rows the
// slab neighbours need); GPF_PART_INTERIOR the rest -- so that a slab driver can exchange the edge
// rows while the interior is being computed.  Grids of fewer than 3 tile rows are not split
// (EDGE = everything, INTERIOR = nothing).
// PART_ROWS (internal): tile rows [row0, row0 + nrows), records from slot rec0, at most max_cta CTAs --
// the row chunks of the pipelined host-buffer step (FvSolver::update_host_pipelined)
constexpr int PART_ROWS = 100;
template <bool CORR>
static int launch_pass(FvSolver& S, int until_converged, cudaStream_t s, int part = GPF_PART_ALL, int row0 = 0,
                       int nrows = 0, int rec0 = 0, int max_cta = 0, int* used_cta = nullptr, SlabComm* xc = nullptr) {
    const int nty = S.P.nblk_x, nbx = S.P.nblk_y;          // tiles along y, tile rows along x
    if (part != PART_ROWS && (part < GPF_PART_ALL || part > GPF_PART_INTERIOR)) { set_error("bad pass part"); return GPF_E_INVALID; }
    if (part != PART_ROWS && nbx < 3) {
        if (part == GPF_PART_INTERIOR) return GPF_OK;
        part = GPF_PART_ALL;
    }
    PassArgs A;
    for (int k = 0; k < 3; ++k) { A.qmap[k] = S.qmap[k]; A.q0map[k] = S.q0map[k]; A.qbuf[k] = S.qbuf[k]; }
    A.topo_map = S.topo_map; A.topo2_map = S.topo2_map; A.ls_map = S.ls_map;
    const bool vf = S.P.visc_field != 0;
    if (vf) {
        // the viscosity of every cell of the pass's input (stress.py:305-326), once per pass: the
        // INTERIOR launch of a split pass reads the plane its EDGE launch left
        A.eta_map = S.eta_map;
        if (part != GPF_PART_INTERIOR) {
            const int rc = S.launch_viscosity(CORR ? 1 : 0, s);
            if (rc) return rc;
        }
    } else {
        A.eta_map = S.ls_map;
    }
    A.P = S.P; A.st = S.st; A.partial = S.partial; A.pflags = S.pflags; A.redo = S.redo;
    A.until_converged = until_converged;
    A.sg_p = S.sg_p; A.sg_xz = S.sg_xz; A.sg_yz = S.sg_yz;
    for (int k = 0; k < 3; ++k) A.sg_scale[k] = S.sg_scale[k];
    memset(&A.X, 0, sizeof(A.X));
    if (xc) {
        // fused exchange (slab_comm.cuh): whole-pass launches only
        if (!xc->attached || part != GPF_PART_ALL) { set_error("fused slab pass needs an attached communicator and a whole-pass launch"); return GPF_E_STATE; }
        SlabX& X = A.X;
        X.on = 1;
        for (int sd = 0; sd < 2; ++sd) {
            for (int k = 0; k < 3; ++k) X.peer_q[sd][k] = xc->peer_q[sd][k];
            X.peer_plane[sd] = xc->peer_plane[sd];
            X.peer_row[sd] = xc->peer_row[sd];
            // my row 1 completes the lower neighbour's upper ghost row (its side 1), my row Nx the upper neighbour's side 0
            X.peer_flag[sd] = xc->nbr[sd] >= 0 ? xc->hflag2(xc->nbr[sd], sd == 0 ? 1 : 0) : nullptr;
            X.edge_dup[sd] = S.P.bc.edge[sd];
        }
        X.my_flag = xc->hflag2(xc->lay.rank, 0);
        X.xseq = xc->xseq(); X.xdone = xc->xdone(); X.err = xc->err();
        X.peers = xc->d_peer; X.off_recs = xc->lay.off_recs; X.off_rflag = xc->lay.off_rflag;
        X.world = xc->lay.world; X.rank = xc->lay.rank; X.rseq = xc->seq(1);
    }
    // INTERIOR leaves a few CTA slots empty so that the exchange kernels of the side stream (pack,
    // send, receive) find room beside the persistent CTAs (GPF_INTERIOR_SLACK, default 2)
    static const int slack = [] { const char* e = getenv("GPF_INTERIOR_SLACK"); return e ? atoi(e) : 2; }();
    const int n_edge = pass_ctas(2 * nty, CORR, nty);
    const int n_int = pass_ctas((nbx - 2) * nty, CORR, nty, slack);
    int n_rows = 0;
    if (part == PART_ROWS) {
        A.ntiles = nrows * nty; A.bx_mul = 1; A.bx_off = row0; A.rec0 = rec0;
        n_rows = std::min(pass_ctas(A.ntiles, CORR, nty), max_cta);
        if (used_cta) *used_cta = n_rows;
        if (CORR) S.nrec_corr = rec0 + n_rows;
    } else if (part == GPF_PART_ALL) {
        A.ntiles = nbx * nty; A.bx_mul = 1; A.bx_off = 0; A.rec0 = 0;
        if (CORR) S.nrec_corr = pass_ctas(A.ntiles, true, nty);
        if (CORR) A.X.nrec = S.nrec_corr;
    } else if (part == GPF_PART_EDGE) {
        A.ntiles = 2 * nty; A.bx_mul = nbx - 1; A.bx_off = 0; A.rec0 = 0;
        if (CORR) S.nrec_corr = n_edge + n_int;
    } else {
        A.ntiles = (nbx - 2) * nty; A.bx_mul = 1; A.bx_off = 1; A.rec0 = n_edge;
        if (CORR) S.nrec_corr = n_edge + n_int;
    }
    const bool ls = S.P.has_extra != 0;
    const bool dh = S.P.eos.kind == GPF_EOS_DH;
    const bool xf = S.P.xflow != 0;
    const int n = (part == PART_ROWS) ? n_rows : (part == GPF_PART_INTERIOR) ? n_int : pass_ctas(A.ntiles, CORR, nty);
    const int rc = (S.cfg.arith == GPF_ARITH_FAST) ? launch_pass_fast(A, CORR, ls, dh, xf, vf, n, s)
                                                   : launch_pass_exact(A, CORR, ls, dh, xf, vf, n, s);
    if (rc) return rc;
    // straight-line kernel + general kernel (returns at once unless flagged); surrogate variant: the general kernel alone
    g_launches += (S.sg_p || S.sg_xz || S.sg_yz) ? 1 : 2;
    return debug_check(CORR ? "fv_pass_kernel<corrector>" : "fv_pass_kernel<predictor>", s);
}

static int launch_finish(FvSolver& S, int nblk, int mode, int until_converged, cudaStream_t s,
                         const Rec5* recs = nullptr, int nrec = 0, Rec5* rec_out = nullptr, int xslot = -1) {
    fv_finish_kernel<<<1, 256, 0, s>>>(S.P, S.st, S.partial, S.pflags, nblk, mode, until_converged, S.refresh,
                                       recs, nrec, rec_out, xslot);
    g_launches++;
    return debug_check("fv_finish_kernel", s);
}

static int launch_reduce(FvSolver& S, cudaStream_t s) {
    if (S.P.eos.kind == GPF_EOS_DH)
        fv_reduce_kernel<GPF_EOS_DH><<<REDUCE_BLOCKS, 256, 0, s>>>(S.P, S.st, S.qtable, S.topo, S.partial, S.pflags);
    else
        fv_reduce_kernel<-1><<<REDUCE_BLOCKS, 256, 0, s>>>(S.P, S.st, S.qtable, S.topo, S.partial, S.pflags);
    g_launches++;
    GPF_CUDA(cudaGetLastError());
    return GPF_OK;
}

int FvSolver::step(int n, int until_converged, cudaStream_t s) {
    if (n < 0) { set_error("n < 0"); return GPF_E_INVALID; }
    if (sg_p || sg_xz || sg_yz) {
        set_error("gpf_fv_step: surrogate closures are set; the GP means must be refreshed before every pass "
                  "(drive gpf_fv_pass / gpf_fv_reduce_local / gpf_fv_finish_surrogate)");
        return GPF_E_INVALID;
    }
    const int nblk = P.ncta;               // = nrec_corr of an unsplit corrector launch
    if (timing) {
        while ((int)ev.size() < 3 * n + 1) { cudaEvent_t e; GPF_CUDA(cudaEventCreate(&e)); ev.push_back(e); }
        ev_steps = n;
        GPF_CUDA(cudaEventRecord(ev[0], s));
    }
    // One step = 5 launches whose arguments never change (all loop state lives in DevState): replayed
    // as a CUDA graph, which takes the launch gaps between them off the step (GPF_NO_GRAPH=1: plain launches)
    static const bool graphs = getenv("GPF_NO_GRAPH") == nullptr && getenv("GPF_DEBUG_SYNC") == nullptr;
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    GPF_CUDA(cudaStreamIsCapturing(s, &cap));
    if (graphs && !timing && n >= 1 && cap == cudaStreamCaptureStatusNone) {
        cudaGraphExec_t& ge = step_graph[until_converged ? 1 : 0];
        if (!ge) {
            // first use of the variant outside a capture: sets the kernels' shared-memory attribute
            int rc;
            if ((rc = launch_pass<false>(*this, until_converged, s))) return rc;
            if ((rc = launch_pass<true>(*this, until_converged, s))) return rc;
            if ((rc = launch_finish(*this, nblk, FIN_STEP, until_converged, s))) return rc;
            --n;
            const long long l0 = g_launches;
            cudaGraph_t g = nullptr;
            cudaStream_t cs;         // the caller's stream may be the legacy default stream, which cannot capture
            GPF_CUDA(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
            cudaError_t e = cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal);
            if (e != cudaSuccess) { cudaStreamDestroy(cs); set_error(std::string("cudaStreamBeginCapture: ") + cudaGetErrorString(e)); return GPF_E_CUDA; }
            rc = launch_pass<false>(*this, until_converged, cs);
            if (!rc) rc = launch_pass<true>(*this, until_converged, cs);
            if (!rc) rc = launch_finish(*this, nblk, FIN_STEP, until_converged, cs);
            e = cudaStreamEndCapture(cs, &g);
            cudaStreamDestroy(cs);
            g_launches = l0;
            if (rc || e != cudaSuccess) {
                if (g) cudaGraphDestroy(g);
                if (!rc) { set_error(std::string("CUDA graph capture of the step failed: ") + cudaGetErrorString(e)); rc = GPF_E_CUDA; }
                return rc;
            }
            e = cudaGraphInstantiate(&ge, g, 0);
            cudaGraphDestroy(g);
            if (e != cudaSuccess) { ge = nullptr; set_error(std::string("cudaGraphInstantiate: ") + cudaGetErrorString(e)); return GPF_E_CUDA; }
        }
        for (int it = 0; it < n; ++it) GPF_CUDA(cudaGraphLaunch(ge, s));
        g_launches += (P.visc_field ? 7LL : 5LL) * n;
        dirty_scalars = false;
        return GPF_OK;
    }
    for (int it = 0; it < n; ++it) {
        int rc;
        if ((rc = launch_pass<false>(*this, until_converged, s))) return rc;
        if (timing) GPF_CUDA(cudaEventRecord(ev[3 * it + 1], s));
        if ((rc = launch_pass<true>(*this, until_converged, s))) return rc;
        if (timing) GPF_CUDA(cudaEventRecord(ev[3 * it + 2], s));
        if ((rc = launch_finish(*this, nblk, FIN_STEP, until_converged, s))) return rc;
        if (timing) GPF_CUDA(cudaEventRecord(ev[3 * it + 3], s));
    }
    dirty_scalars = false;     // the corrector's reductions describe the new q
    return GPF_OK;
}

// ---- one step on a host-resident field, copies overlapped with the passes -------------------
// upload -> predictor -> corrector -> download of a whole field leaves the PCIe link idle in one direction
// at any time and the GPU idle during both copies.  Here the field moves in chunks of tile rows on two
// copy streams while the pass kernels follow chunk by chunk on the caller's stream.  For the MacCormack
// order predictor d = +1 / corrector d = -1 (switch = +1, problem.py:521-522) the dependencies are
//   P(k) reads input rows [a_k - 1, b_k]                  -> after upload(0..k)
//   C(k) reads predictor rows [a_k, b_k + 1], q0 [a_k, b_k] -> after P(k + 1)   (the last chunk: the ghost row
//        Nx+1, written by the predictor's own boundary tiles: P(0) for a periodic x side, P(K-1) otherwise)
//   download(k)                                            -> after C(k); the two x ghost rows after the last C
// so the launch order P(0) P(1) C(0) P(2) C(1) ... C(K-1) on one stream satisfies all of them, and a chunk
// launch is the same kernel with a tile-row range: the result is bit-identical to gpf_fv_step.
// Returns GPF_E_STATE when the configuration is not eligible (the caller then takes the plain path).
int FvSolver::update_host_pipelined(double* q_host, cudaStream_t s) {
    const int nbx = P.nblk_y;
    if (P.mc_order != 1 || P.visc_field || sg_p || sg_xz || sg_yz || timing || nbx < 8 ||
        P.bc.type[0][0] == GPF_BC_HALO || P.bc.type[1][0] == GPF_BC_HALO)
        return GPF_E_STATE;
    static const int want = [] { const char* e = getenv("GPF_HOST_CHUNKS"); return e ? atoi(e) : 16; }();
    if (want < 2) return GPF_E_STATE;
    const int K = std::min(want, nbx / 4);
    int cur, rc;
    if ((rc = read_cur(s, &cur))) return rc;                 // synchronises s
    if (!cp_stream[0]) {
        for (int k = 0; k < 2; ++k) GPF_CUDA(cudaStreamCreateWithFlags(&cp_stream[k], cudaStreamNonBlocking));
        cp_event.resize(2 * 64 + 2);
        for (auto& e : cp_event) GPF_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    if (K > 64) return GPF_E_STATE;
    const size_t Ny2 = cfg.Ny + 2, Nx2 = cfg.Nx + 2, pitch = P.L.pitch;
    double* qin = qbuf[cur];
    const double* qout = qbuf[(cur + 2) % 3];
    // chunk k: tile rows [t0, t1); padded-array rows uploaded with it: the interior rows of its tiles, chunk 0
    // also ghost row 0, the last chunk also ghost row Nx+1
    auto trow = [&](int k) { return (int)((long long)nbx * k / K); };
    auto rows = [&](int k, size_t* r0, size_t* r1, bool with_ghost) {
        *r0 = (size_t)trow(k) * TX + 1;
        *r1 = std::min<size_t>((size_t)trow(k + 1) * TX + 1, (size_t)cfg.Nx + 1);
        if (with_ghost && k == 0) *r0 = 0;
        if (with_ghost && k == K - 1) *r1 = Nx2;
    };
    cudaEvent_t ev_start = cp_event[2 * 64];
    GPF_CUDA(cudaEventRecord(ev_start, s));
    GPF_CUDA(cudaStreamWaitEvent(cp_stream[0], ev_start, 0));
    GPF_CUDA(cudaStreamWaitEvent(cp_stream[1], ev_start, 0));
    for (int k = 0; k < K; ++k) {
        size_t r0, r1;
        rows(k, &r0, &r1, true);
        for (int c = 0; c < 3; ++c)
            GPF_CUDA(cudaMemcpy2DAsync(qin + c * P.L.plane + r0 * pitch + 1, pitch * 8, q_host + (size_t)c * Nx2 * Ny2 + r0 * Ny2,
                                       Ny2 * 8, Ny2 * 8, r1 - r0, cudaMemcpyHostToDevice, cp_stream[0]));
        GPF_CUDA(cudaEventRecord(cp_event[k], cp_stream[0]));
    }
    const int per = std::max(1, (npartial - 1) / K);         // record slots of one corrector chunk
    int rec0 = 0;
    auto corr_chunk = [&](int k) -> int {
        int used = 0;
        int r = launch_pass<true>(*this, 0, s, PART_ROWS, trow(k), trow(k + 1) - trow(k), rec0, per, &used);
        if (r) return r;
        rec0 += used;
        GPF_CUDA(cudaEventRecord(cp_event[64 + k], s));
        GPF_CUDA(cudaStreamWaitEvent(cp_stream[1], cp_event[64 + k], 0));
        size_t r0, r1;
        rows(k, &r0, &r1, false);
        for (int c = 0; c < 3; ++c)
            GPF_CUDA(cudaMemcpy2DAsync(q_host + (size_t)c * Nx2 * Ny2 + r0 * Ny2, Ny2 * 8, qout + c * P.L.plane + r0 * pitch + 1,
                                       pitch * 8, Ny2 * 8, r1 - r0, cudaMemcpyDeviceToHost, cp_stream[1]));
        return GPF_OK;
    };
    for (int k = 0; k < K; ++k) {
        GPF_CUDA(cudaStreamWaitEvent(s, cp_event[k], 0));
        if ((rc = launch_pass<false>(*this, 0, s, PART_ROWS, trow(k), trow(k + 1) - trow(k), 0, 1 << 30))) return rc;
        if (k >= 1 && (rc = corr_chunk(k - 1))) return rc;
    }
    if ((rc = corr_chunk(K - 1))) return rc;
    // the two x ghost rows of the new field (written by the first and the last chunk's boundary tiles)
    cudaEvent_t ev_c = cp_event[2 * 64 + 1];
    GPF_CUDA(cudaEventRecord(ev_c, s));
    GPF_CUDA(cudaStreamWaitEvent(cp_stream[1], ev_c, 0));
    for (int c = 0; c < 3; ++c)
        for (size_t r : {(size_t)0, Nx2 - 1})
            GPF_CUDA(cudaMemcpyAsync(q_host + (size_t)c * Nx2 * Ny2 + r * Ny2, qout + c * P.L.plane + r * pitch + 1, Ny2 * 8,
                                     cudaMemcpyDeviceToHost, cp_stream[1]));
    if ((rc = launch_finish(*this, nrec_corr, FIN_STEP, 0, s))) return rc;
    dirty_scalars = false;
    refresh_valid = false;
    GPF_CUDA(cudaEventRecord(ev_start, cp_stream[1]));
    GPF_CUDA(cudaStreamWaitEvent(s, ev_start, 0));           // the caller's stream owns the completed download
    return GPF_OK;
}

// ---- slab decomposition building blocks (gapflow_b200/dist.py drives them) ---------------
int FvSolver::pass(int corrector, int part, cudaStream_t s) {
    return corrector ? launch_pass<true>(*this, run_mode_uc, s, part) : launch_pass<false>(*this, run_mode_uc, s, part);
}

int FvSolver::halo_pack(int rel, double* send, cudaStream_t s) {
    if (rel < 0 || rel > 2) { set_error("rel must be 0, 1 or 2"); return GPF_E_INVALID; }
    const int n = 6 * (cfg.Ny + 2);
    fv_halo_pack_kernel<<<(n + 255) / 256, 256, 0, s>>>(P, st, qtable, rel, send);
    g_launches++;
    return debug_check("fv_halo_pack_kernel", s);
}

int FvSolver::halo_unpack(int rel, const double* recv, cudaStream_t s) {
    if (rel < 0 || rel > 2) { set_error("rel must be 0, 1 or 2"); return GPF_E_INVALID; }
    const int n = 6 * (cfg.Ny + 2);
    fv_halo_unpack_kernel<<<(n + 255) / 256, 256, 0, s>>>(P, st, qtable, rel, recv);
    g_launches++;
    if (rel == 0) { dirty_scalars = true; refresh_valid = false; }
    int rc = debug_check("fv_halo_unpack_kernel", s);
    if (rc) return rc;
    if (rel == 2 && (P.bc.edge[0] || P.bc.edge[1])) {
        const int slot = npartial - 1;       // fixed slot, folded as the extra record by reduce_local
        if (P.eos.kind == GPF_EOS_DH) fv_edge_rows_kernel<GPF_EOS_DH><<<1, 256, 0, s>>>(P, st, qtable, rel, partial, pflags, slot);
        else fv_edge_rows_kernel<-1><<<1, 256, 0, s>>>(P, st, qtable, rel, partial, pflags, slot);
        g_launches++;
        rc = debug_check("fv_edge_rows_kernel", s);
    }
    return rc;
}

// kind 0: fold the records of the last corrector pass; kind 1: reduce the current q first
int FvSolver::reduce_local(int kind, double* rec5, cudaStream_t s) {
    int rc;
    int nblk = nrec_corr;
    int xslot = (P.bc.edge[0] || P.bc.edge[1]) ? npartial - 1 : -1;
    if (kind == 1) {
        if ((rc = launch_reduce(*this, s))) return rc;
        nblk = REDUCE_BLOCKS;
        xslot = -1;
    }
    return launch_finish(*this, nblk, FIN_REFRESH, 0, s, nullptr, 0, reinterpret_cast<Rec5*>(rec5), xslot);
}

int FvSolver::finish(int mode, const double* recs, int nrec, cudaStream_t s) {
    if (mode < FIN_STEP || mode > FIN_REFRESH || !recs || nrec < 1) { set_error("bad finish arguments"); return GPF_E_INVALID; }
    int rc = launch_finish(*this, 0, mode, run_mode_uc, s, reinterpret_cast<const Rec5*>(recs), nrec);
    if (rc) return rc;
    if (mode != FIN_REFRESH) dirty_scalars = false;
    else refresh_valid = true;
    return GPF_OK;
}

// ---- GP surrogate closures (stress.py:346-356, 614-620) -------------------------------------
int FvSolver::set_surrogate(const double* p, double ps, const double* xz, double xzs, const double* yz, double yzs) {
    if ((p || xz || yz) && P.visc_field) {
        set_error("gpf_fv_set_surrogate: a solver with viscosity modifiers has no surrogate variant");
        return GPF_E_INVALID;
    }
    sg_p = p; sg_xz = xz; sg_yz = yz;
    sg_scale[0] = ps; sg_scale[1] = xzs; sg_scale[2] = yzs;
    return GPF_OK;
}

// rec5[2] (max c^2) <- c2 when the pressure is a GP (Pressure.v_sound, stress.py:522-537); the flag of
// an undefined analytic sound speed is replaced by that of c2
__global__ void fv_record_c2_kernel(double* recs, int nrec, double c2) {
    const int r = threadIdx.x;
    if (r >= nrec) return;
    recs[r * 5 + 2] = c2;
    int f = (int)recs[r * 5 + 4] & ~FLAG_CNAN;
    if (!(c2 >= 0.0)) f |= FLAG_CNAN;       // sqrt of a negative slope is NaN (stress.py:535)
    recs[r * 5 + 4] = (double)f;
}

int FvSolver::finish_surrogate(int mode, double* recs, int nrec, double c2, cudaStream_t s) {
    if (!recs || nrec < 1 || nrec > 1024) { set_error("bad record arguments"); return GPF_E_INVALID; }
    if (sg_p) {
        fv_record_c2_kernel<<<1, 1024, 0, s>>>(recs, nrec, c2);
        g_launches++;
        GPF_CUDA(cudaGetLastError());
    }
    return finish(mode, recs, nrec, s);
}

// ---- peer-memory exchange (slab_comm.cuh) --------------------------------------------------
SlabCommLayout SlabComm::layout(int world, int rank, int Ny2) {
    SlabCommLayout l;
    l.world = world; l.rank = rank; l.Ny2 = Ny2;
    size_t o = 0;
    l.off_halo = o; o += align_up((size_t)2 * 2 * 3 * Ny2 * 8, 256);
    l.off_hflag = o; o += 256;
    l.off_recs = o; o += align_up((size_t)2 * world * 8 * 8, 256);
    l.off_rflag = o; o += align_up((size_t)world * 8, 256);
    l.off_err = o; o += 256;
    l.off_seq = o; o += 256;
    l.off_ptrs = o; o += align_up((size_t)world * sizeof(void*), 256);
    l.off_x = o; o += 256;
    l.bytes = o;
    return l;
}

int SlabComm::create(int world, int rank, int Ny) {
    if (world < 1 || rank < 0 || rank >= world || Ny < 1 || world > 256) { set_error("bad communicator shape"); return GPF_E_INVALID; }
    lay = layout(world, rank, Ny + 2);
    GPF_CUDA(cudaMalloc(&local, lay.bytes));
    GPF_CUDA(cudaMemset(local, 0, lay.bytes));
    {
        const char* e = getenv("GPF_COMM_TIMEOUT_S");
        const double sec = (e && atof(e) > 0.0) ? atof(e) : 30.0;
        int rc = set_timeout(sec);
        if (rc) return rc;
    }
    GPF_CUDA(cudaDeviceSynchronize());
    GPF_CUDA(cudaIpcGetMemHandle(&handle, local));
    peer.assign(world, nullptr);
    peer[rank] = local;
    d_peer = reinterpret_cast<unsigned char**>(local + lay.off_ptrs);
    return GPF_OK;
}

int SlabComm::set_timeout(double seconds) {
    if (!local || !(seconds > 0.0) || seconds > 3600.0) { set_error("gpf_comm_set_timeout: 0 < seconds <= 3600"); return GPF_E_INVALID; }
    const unsigned long long ns = (unsigned long long)(seconds * 1e9);
    GPF_CUDA(cudaMemcpy(local + lay.off_err + 8, &ns, sizeof(ns), cudaMemcpyHostToDevice));
    return GPF_OK;
}

int SlabComm::connect(const unsigned char* handles) {
    for (int r = 0; r < lay.world; ++r) {
        if (r == lay.rank) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, handles + (size_t)r * sizeof(h), sizeof(h));
        void* p = nullptr;
        GPF_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        peer[r] = static_cast<unsigned char*>(p);
    }
    GPF_CUDA(cudaMemcpy(d_peer, peer.data(), (size_t)lay.world * sizeof(void*), cudaMemcpyHostToDevice));
    connected = true;
    return GPF_OK;
}

// ---- fused exchange: the neighbours' q buffers ---------------------------------------------------------
struct PeerBlob {                      // what a rank tells the others about its rotating q buffers
    cudaIpcMemHandle_t handle;         // of the allocation that holds the solver's workspace
    unsigned long long offset;         // of qbuf[0] inside it (bytes)
    unsigned long long plane;          // plane stride (doubles); buffer k starts 3 k planes on
    int pitch, Nx, Ny, pad;
};
static_assert(sizeof(PeerBlob) == GPF_PEER_BLOB_BYTES, "peer blob size");

typedef CUresult (*AddrRangeFn)(CUdeviceptr*, size_t*, CUdeviceptr);

int FvSolver::peer_export(unsigned char* blob) {
    static AddrRangeFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &p, cudaEnableDefault, &qres) != cudaSuccess || qres != cudaDriverEntryPointSuccess) {
            cudaGetLastError();
            set_error("cuMemGetAddressRange not available from the driver"); return GPF_E_CUDA;
        }
        fn = reinterpret_cast<AddrRangeFn>(p);
    }
    CUdeviceptr base = 0; size_t size = 0;
    if (fn(&base, &size, reinterpret_cast<CUdeviceptr>(qbuf[0])) != CUDA_SUCCESS) { set_error("cuMemGetAddressRange failed on the workspace"); return GPF_E_CUDA; }
    PeerBlob b;
    memset(&b, 0, sizeof(b));
    // a workspace from a stream-ordered or virtual-memory allocator has no legacy IPC handle: the caller then keeps
    // the exchange kernels (gpf_comm_attach reports it)
    cudaError_t e = cudaIpcGetMemHandle(&b.handle, reinterpret_cast<void*>(base));
    if (e != cudaSuccess) { cudaGetLastError(); return cuda_fail(e, "cudaIpcGetMemHandle(workspace)"); }
    b.offset = reinterpret_cast<unsigned long long>(qbuf[0]) - (unsigned long long)base;
    b.plane = P.L.plane; b.pitch = P.L.pitch; b.Nx = cfg.Nx; b.Ny = cfg.Ny;
    memcpy(blob, &b, sizeof(b));
    return GPF_OK;
}

static int attach_checks(const FvSolver& S, const SlabComm& c, int lo, int hi) {
    if (!c.connected) { set_error("communicator not connected"); return GPF_E_STATE; }
    if (lo >= c.lay.world || hi >= c.lay.world) { set_error("bad neighbour rank"); return GPF_E_INVALID; }
    if ((lo >= 0) != (S.P.bc.type[0][0] == GPF_BC_HALO) || (hi >= 0) != (S.P.bc.type[1][0] == GPF_BC_HALO)) {
        set_error("gpf_comm_attach: neighbours do not match the solver's HALO sides"); return GPF_E_INVALID;
    }
    return GPF_OK;
}

int SlabComm::attach(const FvSolver& S, const unsigned char* blobs, int lo, int hi) {
    int rc = attach_checks(S, *this, lo, hi);
    if (rc) return rc;
    const int nb[2] = {lo, hi};
    for (int sd = 0; sd < 2; ++sd) {
        nbr[sd] = nb[sd];
        if (nb[sd] < 0) continue;
        PeerBlob b;
        memcpy(&b, blobs + (size_t)nb[sd] * sizeof(b), sizeof(b));
        if (b.pitch != S.P.L.pitch || b.Ny != S.cfg.Ny) { set_error("gpf_comm_attach: the neighbour's rows have another length"); return GPF_E_INVALID; }
        void* p = nullptr;
        if (sd == 1 && nb[1] == nb[0]) p = peer_map[0];           // two ranks on a ring: one peer on both sides
        else GPF_CUDA(cudaIpcOpenMemHandle(&p, b.handle, cudaIpcMemLazyEnablePeerAccess));
        peer_map[sd] = p;
        double* q0 = reinterpret_cast<double*>(static_cast<unsigned char*>(p) + b.offset);
        for (int k = 0; k < 3; ++k) peer_q[sd][k] = q0 + (size_t)k * 3 * b.plane;
        peer_plane[sd] = b.plane;
        peer_row[sd] = (sd == 0) ? b.Nx + 1 : 0;
    }
    attached = true;
    return GPF_OK;
}

int SlabComm::attach_local(const FvSolver& S, const FvSolver* lo_s, int lo, const FvSolver* hi_s, int hi) {
    int rc = attach_checks(S, *this, lo, hi);
    if (rc) return rc;
    const FvSolver* ns[2] = {lo_s, hi_s};
    const int nb[2] = {lo, hi};
    for (int sd = 0; sd < 2; ++sd) {
        nbr[sd] = nb[sd];
        if (nb[sd] < 0) continue;
        if (!ns[sd] || ns[sd]->P.L.pitch != S.P.L.pitch) { set_error("gpf_comm_attach_local: bad neighbour solver"); return GPF_E_INVALID; }
        for (int k = 0; k < 3; ++k) peer_q[sd][k] = ns[sd]->qbuf[k];
        peer_plane[sd] = ns[sd]->P.L.plane;
        peer_row[sd] = (sd == 0) ? ns[sd]->cfg.Nx + 1 : 0;
    }
    attached = true;
    return GPF_OK;
}

int SlabComm::detach() {
    if (!local_only) {
        if (peer_map[0]) cudaIpcCloseMemHandle(peer_map[0]);
        if (peer_map[1] && peer_map[1] != peer_map[0]) cudaIpcCloseMemHandle(peer_map[1]);
    }
    cudaGetLastError();
    peer_map[0] = peer_map[1] = nullptr;
    for (int sd = 0; sd < 2; ++sd)
        for (int k = 0; k < 3; ++k) peer_q[sd][k] = nullptr;
    attached = false;
    if (step_graph) { cudaGraphExecDestroy(step_graph); step_graph = nullptr; }     // captured with the fused passes
    return GPF_OK;
}

int SlabComm::connect_local(SlabComm* const* all) {
    for (int r = 0; r < lay.world; ++r) {
        if (!all[r] || !all[r]->local || all[r]->lay.world != lay.world || all[r]->lay.Ny2 != lay.Ny2) { set_error("gpf_comm_connect_local: bad communicator table"); return GPF_E_INVALID; }
        peer[r] = all[r]->local;
    }
    GPF_CUDA(cudaMemcpy(d_peer, peer.data(), (size_t)lay.world * sizeof(void*), cudaMemcpyHostToDevice));
    connected = true;
    local_only = true;
    return GPF_OK;
}

int SlabComm::destroy() {
    if (step_graph) { cudaGraphExecDestroy(step_graph); step_graph = nullptr; }
    if (!local_only) {
        if (peer_map[0]) cudaIpcCloseMemHandle(peer_map[0]);
        if (peer_map[1] && peer_map[1] != peer_map[0]) cudaIpcCloseMemHandle(peer_map[1]);
    }
    peer_map[0] = peer_map[1] = nullptr;
    if (local_only) peer.clear();
    for (int r = 0; r < (int)peer.size(); ++r)
        if (r != lay.rank && peer[r]) cudaIpcCloseMemHandle(peer[r]);
    peer.clear();
    if (local) cudaFree(local);
    local = nullptr;
    return GPF_OK;
}

static int comm_ok(const FvSolver& S, const SlabComm& c) {
    if (!c.connected) { set_error("communicator not connected"); return GPF_E_STATE; }
    if (c.lay.Ny2 != S.cfg.Ny + 2) { set_error("communicator built for another Ny"); return GPF_E_INVALID; }
    return GPF_OK;
}

int FvSolver::halo_send(SlabComm& c, int rel, int lo, int hi, cudaStream_t s) {
    int rc = comm_ok(*this, c);
    if (rc) return rc;
    if (rel < 0 || rel > 2 || lo >= c.lay.world || hi >= c.lay.world) { set_error("bad halo_send arguments"); return GPF_E_INVALID; }
    // my row 1 is the lower neighbour's upper ghost row (its side 1); my row Nx the upper neighbour's side 0
    const int nblk = std::max(1, std::min(128, (6 * (cfg.Ny + 2) + 255) / 256));
    fv_halo_send_kernel<<<nblk, 256, 0, s>>>(P, st, qtable, rel, lo >= 0 ? c.halo(lo, 0, 1) : nullptr,
                                             hi >= 0 ? c.halo(hi, 0, 0) : nullptr, c.halo_par_stride(),
                                             lo >= 0 ? c.hflag(lo, 1) : nullptr, hi >= 0 ? c.hflag(hi, 0) : nullptr, c.seq(0),
                                             c.done(0));
    g_launches++;
    return debug_check("fv_halo_send_kernel", s);
}

int FvSolver::halo_recv(SlabComm& c, int rel, cudaStream_t s) {
    int rc = comm_ok(*this, c);
    if (rc) return rc;
    if (rel < 0 || rel > 2) { set_error("rel must be 0, 1 or 2"); return GPF_E_INVALID; }
    const int nblk = std::max(1, std::min(128, (6 * (cfg.Ny + 2) + 255) / 256));
    fv_halo_recv_kernel<<<nblk, 256, 0, s>>>(P, st, qtable, rel, c.halo(c.lay.rank, 0, 0), c.halo_par_stride(),
                                             c.hflag(c.lay.rank, 0), c.seq(0), c.err(), c.done(1));
    g_launches++;
    if (rel == 0) { dirty_scalars = true; refresh_valid = false; }
    rc = debug_check("fv_halo_recv_kernel", s);
    if (rc) return rc;
    if (rel == 2 && (P.bc.edge[0] || P.bc.edge[1])) {
        const int slot = npartial - 1;
        if (P.eos.kind == GPF_EOS_DH) fv_edge_rows_kernel<GPF_EOS_DH><<<1, 256, 0, s>>>(P, st, qtable, rel, partial, pflags, slot);
        else fv_edge_rows_kernel<-1><<<1, 256, 0, s>>>(P, st, qtable, rel, partial, pflags, slot);
        g_launches++;
        rc = debug_check("fv_edge_rows_kernel", s);
    }
    return rc;
}

int FvSolver::record_send(SlabComm& c, int kind, cudaStream_t s) {
    int rc = comm_ok(*this, c);
    if (rc) return rc;
    int nblk = nrec_corr;
    int xslot = (P.bc.edge[0] || P.bc.edge[1]) ? npartial - 1 : -1;
    if (kind == 1) {
        if ((rc = launch_reduce(*this, s))) return rc;
        nblk = REDUCE_BLOCKS;
        xslot = -1;
    }
    fv_record_send_kernel<<<1, 256, 0, s>>>(partial, pflags, nblk, xslot, c.d_peer, c.lay.off_recs, c.lay.off_rflag,
                                            c.lay.world, c.lay.rank, c.seq(1), c.err());
    g_launches++;
    return debug_check("fv_record_send_kernel", s);
}

int FvSolver::finish_recv(SlabComm& c, int mode, double* recs_out, cudaStream_t s, int until_converged) {
    int rc = comm_ok(*this, c);
    if (rc) return rc;
    if (mode < FIN_STEP || mode > FIN_REFRESH) { set_error("bad finish mode"); return GPF_E_INVALID; }
    fv_finish_kernel<<<1, 256, 0, s>>>(P, st, partial, pflags, 0, mode, until_converged, refresh,
                                       reinterpret_cast<const Rec5*>(c.recs(c.lay.rank, 0)), c.lay.world, nullptr, -1,
                                       c.rflag(c.lay.rank), c.seq(1), c.err(), 8, recs_out);
    g_launches++;
    if ((rc = debug_check("fv_finish_kernel", s))) return rc;
    if (mode != FIN_REFRESH) dirty_scalars = false;
    else refresh_valid = true;
    return GPF_OK;
}

// n x Problem.update() of this rank's slab: pass, edge rows to the neighbours, ghost rows from them
// (twice), records to every rank, fold.  All sequence numbers live on the device, so one step is
// captured once as a CUDA graph (10 kernels) and replayed.
int FvSolver::slab_pass(SlabComm& c, int corrector, cudaStream_t s) {
    int rc = comm_ok(*this, c);
    if (rc) return rc;
    return corrector ? launch_pass<true>(*this, run_mode_uc, s, GPF_PART_ALL, 0, 0, 0, 0, nullptr, &c)
                     : launch_pass<false>(*this, run_mode_uc, s, GPF_PART_ALL, 0, 0, 0, 0, nullptr, &c);
}

int FvSolver::slab_step_once(SlabComm& c, int lo, int hi, double* recs_out, cudaStream_t s) {
    int rc;
    // fused: the boundary tiles of the pass kernels store the edge rows into the neighbours' ghost rows, the general
    // launch of each pass raises the flags and (corrector) publishes the folded record: 5 kernels per step like a
    // single domain, none of them an exchange kernel (GPF_SLAB_FUSED=0: the send / receive / record kernels below)
    static const bool fused = [] { const char* e = getenv("GPF_SLAB_FUSED"); return !e || atoi(e) != 0; }();
    if (fused && c.attached && c.nbr[0] == lo && c.nbr[1] == hi && !P.visc_field && !(sg_p || sg_xz || sg_yz)) {
        if ((rc = slab_pass(c, 0, s))) return rc;
        if ((rc = slab_pass(c, 1, s))) return rc;
        return finish_recv(c, FIN_STEP, recs_out, s, run_mode_uc);
    }
    if ((rc = launch_pass<false>(*this, run_mode_uc, s))) return rc;
    if ((rc = halo_send(c, 1, lo, hi, s))) return rc;
    if ((rc = halo_recv(c, 1, s))) return rc;
    if ((rc = launch_pass<true>(*this, run_mode_uc, s))) return rc;
    if ((rc = halo_send(c, 2, lo, hi, s))) return rc;
    if ((rc = halo_recv(c, 2, s))) return rc;
    if ((rc = record_send(c, 0, s))) return rc;
    return finish_recv(c, FIN_STEP, recs_out, s, run_mode_uc);
}

int FvSolver::slab_step(SlabComm& c, int lo, int hi, int n, double* recs_out, cudaStream_t s) {
    if (n < 0) { set_error("n < 0"); return GPF_E_INVALID; }
    int rc = comm_ok(*this, c);
    if (rc) return rc;
    static const bool graphs = getenv("GPF_NO_GRAPH") == nullptr && getenv("GPF_DEBUG_SYNC") == nullptr;
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    GPF_CUDA(cudaStreamIsCapturing(s, &cap));
    if (!graphs || cap != cudaStreamCaptureStatusNone) {
        for (int it = 0; it < n; ++it)
            if ((rc = slab_step_once(c, lo, hi, recs_out, s))) return rc;
        return GPF_OK;
    }
    if (c.step_graph && (c.graph_lo != lo || c.graph_hi != hi || c.graph_owner != this || c.graph_recs != recs_out || graph_stale)) {
        cudaGraphExecDestroy(c.step_graph);
        c.step_graph = nullptr;
    }
    if (!c.step_graph && n > 0) {
        if ((rc = slab_step_once(c, lo, hi, recs_out, s))) return rc;     // also sets the kernels' attributes outside a capture
        --n;
        const long long l0 = g_launches;
        cudaStream_t cs;
        cudaGraph_t g = nullptr;
        GPF_CUDA(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
        cudaError_t e = cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal);
        if (e != cudaSuccess) { cudaStreamDestroy(cs); set_error(std::string("cudaStreamBeginCapture: ") + cudaGetErrorString(e)); return GPF_E_CUDA; }
        rc = slab_step_once(c, lo, hi, recs_out, cs);
        e = cudaStreamEndCapture(cs, &g);
        cudaStreamDestroy(cs);
        slab_launches = (int)(g_launches - l0);
        g_launches = l0;
        if (rc || e != cudaSuccess) {
            if (g) cudaGraphDestroy(g);
            if (!rc) { set_error(std::string("CUDA graph capture of the slab step failed: ") + cudaGetErrorString(e)); rc = GPF_E_CUDA; }
            return rc;
        }
        e = cudaGraphInstantiate(&c.step_graph, g, 0);
        cudaGraphDestroy(g);
        if (e != cudaSuccess) { c.step_graph = nullptr; set_error(std::string("cudaGraphInstantiate: ") + cudaGetErrorString(e)); return GPF_E_CUDA; }
        c.graph_lo = lo; c.graph_hi = hi; c.graph_owner = this; c.graph_recs = recs_out;
        graph_stale = false;
    }
    for (int it = 0; it < n; ++it) GPF_CUDA(cudaGraphLaunch(c.step_graph, s));
    g_launches += (long long)slab_launches * n;
    dirty_scalars = false;
    return GPF_OK;
}

int FvSolver::get_timing(double* ms_pred, double* ms_corr, long long* launches) {
    double a = 0, b = 0;
    if (timing && ev_steps > 0) {
        GPF_CUDA(cudaEventSynchronize(ev[3 * ev_steps]));
        for (int it = 0; it < ev_steps; ++it) {
            float t;
            GPF_CUDA(cudaEventElapsedTime(&t, ev[3 * it], ev[3 * it + 1])); a += t;
            GPF_CUDA(cudaEventElapsedTime(&t, ev[3 * it + 1], ev[3 * it + 2])); b += t;
        }
        a /= ev_steps; b /= ev_steps;
    }
    if (ms_pred) *ms_pred = a;
    if (ms_corr) *ms_corr = b;
    if (launches) *launches = 3LL * ev_steps;
    return GPF_OK;
}

int FvSolver::pre_run(int reset_ekin_old, cudaStream_t s) {
    int rc;
    if ((rc = launch_reduce(*this, s))) return rc;
    if ((rc = launch_finish(*this, REDUCE_BLOCKS, reset_ekin_old ? FIN_PRERUN : FIN_PRERUN_KEEP_EKIN, 0, s))) return rc;
    GPF_CUDA(cudaStreamSynchronize(s));
    dirty_scalars = false;
    return GPF_OK;
}

int FvSolver::refresh_scalars(cudaStream_t s) {
    int rc;
    if ((rc = launch_reduce(*this, s))) return rc;
    return launch_finish(*this, REDUCE_BLOCKS, FIN_REFRESH, 0, s);
}

int FvSolver::fill_ghosts(cudaStream_t s) {
    const int n = 2 * cfg.Nx + 2 * cfg.Ny;
    fv_fill_ghosts_kernel<<<(n + 127) / 128, 128, 0, s>>>(P, st, qtable);
    g_launches++;
    GPF_CUDA(cudaGetLastError());
    dirty_scalars = true;
    refresh_valid = false;
    return GPF_OK;
}

int FvSolver::get_scalars(gpf_fv_scalars* o, cudaStream_t s) {
    DevState h;
    RefreshOut r;
    if (dirty_scalars) {
        if (!refresh_valid) {      // slab mode: dist.py has already folded all ranks' records (finish(REFRESH))
            int rc = refresh_scalars(s);
            if (rc) return rc;
        }
        GPF_CUDA(cudaMemcpyAsync(&r, refresh, sizeof(r), cudaMemcpyDeviceToHost, s));
    }
    GPF_CUDA(cudaMemcpyAsync(&h, st, sizeof(h), cudaMemcpyDeviceToHost, s));
    GPF_CUDA(cudaStreamSynchronize(s));
    if (h.comm_err) {
        set_error("slab exchange: a wait on a peer rank timed out (a peer died or lagged beyond gpf_comm_set_timeout); "
                  "the run was stopped on every rank and the fields of the last step are not valid");
        return GPF_E_STATE;
    }
    memset(o, 0, sizeof(*o));
    o->step = h.step; o->simtime = h.simtime; o->dt = h.dt; o->residual = h.residual;
    for (int k = 0; k < 5; ++k) o->residual_buffer[k] = h.resbuf[k];
    o->residual_count = h.rescount;
    o->ekin_old = h.ekin_old;
    if (dirty_scalars) {
        o->ekin = r.ekin; o->v_max = r.vmax; o->v_sound = r.vsound;
        const double dxy = (cfg.dy < cfg.dx) ? cfg.dy : cfg.dx;
        o->dt_crit = dxy / (r.vmax + r.vsound);
        o->cfl = h.dt / o->dt_crit;
    } else {
        o->ekin = h.ekin; o->v_max = h.vmax; o->v_sound = h.vsound; o->dt_crit = h.dt_crit;
        o->cfl = h.dt / h.dt_crit;
    }
    o->has_nan = h.has_nan; o->has_neg = h.has_neg; o->stop = h.stop; o->converged = h.converged;
    return GPF_OK;
}

int FvSolver::set_scalars(const gpf_fv_scalars* in, cudaStream_t s) {
    DevState h;
    GPF_CUDA(cudaMemcpyAsync(&h, st, sizeof(h), cudaMemcpyDeviceToHost, s));
    GPF_CUDA(cudaStreamSynchronize(s));
    h.step = in->step; h.simtime = in->simtime; h.dt = in->dt; h.residual = in->residual;
    h.rescount = in->residual_count < 1 ? 0 : (in->residual_count > 5 ? 5 : in->residual_count);
    for (int k = 0; k < 5; ++k) h.resbuf[k] = in->residual_buffer[k];
    h.ekin_old = in->ekin_old;
    h.stop = in->stop;
    int conv = 1;
    for (int k = 0; k < h.rescount; ++k) conv &= (h.resbuf[k] < cfg.tol) ? 1 : 0;
    h.converged = conv;
    h.halt = (conv || h.step >= cfg.max_it || h.stop) ? 1 : 0;
    GPF_CUDA(cudaMemcpyAsync(st, &h, sizeof(h), cudaMemcpyHostToDevice, s));
    GPF_CUDA(cudaStreamSynchronize(s));
    return GPF_OK;
}

int FvSolver::mass(double* m, cudaStream_t s) {
    int rc = refresh_scalars(s);
    if (rc) return rc;
    RefreshOut r;
    GPF_CUDA(cudaMemcpyAsync(&r, refresh, sizeof(r), cudaMemcpyDeviceToHost, s));
    GPF_CUDA(cudaStreamSynchronize(s));
    *m = r.mass;
    return GPF_OK;
}

int FvSolver::closures(double* p, double* wxz, double* wyz, double* bulk, int on_device, cudaStream_t s, int rel) {
    const size_t np = (size_t)(cfg.Nx + 2) * (cfg.Ny + 2);
    double* d[4] = {p, wxz, wyz, bulk};
    const int nc[4] = {1, 12, 12, 3};
    double* tmp[4] = {nullptr, nullptr, nullptr, nullptr};
    if (!on_device)
        for (int k = 0; k < 4; ++k)
            if (d[k]) GPF_CUDA(cudaMallocAsync(&tmp[k], nc[k] * np * 8, s));
    double** o = on_device ? d : tmp;
    const int blocks = (int)((np + 255) / 256);
    const bool ls = P.has_extra != 0, dh = P.eos.kind == GPF_EOS_DH;
    const double* ef = nullptr;          // viscosity field of the same state
    if (P.visc_field) {
        int rc = launch_viscosity(rel, s);
        if (rc) return rc;
        ef = eta;
    }
    if (ls) { if (dh) fv_closures_kernel<true, GPF_EOS_DH><<<blocks, 256, 0, s>>>(P, st, qtable, topo, extra, o[0], o[1], o[2], o[3], rel, ef);
              else    fv_closures_kernel<true, -1><<<blocks, 256, 0, s>>>(P, st, qtable, topo, extra, o[0], o[1], o[2], o[3], rel, ef); }
    else    { if (dh) fv_closures_kernel<false, GPF_EOS_DH><<<blocks, 256, 0, s>>>(P, st, qtable, topo, extra, o[0], o[1], o[2], o[3], rel, ef);
              else    fv_closures_kernel<false, -1><<<blocks, 256, 0, s>>>(P, st, qtable, topo, extra, o[0], o[1], o[2], o[3], rel, ef); }
    g_launches++;
    GPF_CUDA(cudaGetLastError());
    if (!on_device) {
        for (int k = 0; k < 4; ++k)
            if (d[k]) {
                GPF_CUDA(cudaMemcpyAsync(d[k], tmp[k], nc[k] * np * 8, cudaMemcpyDeviceToHost, s));
                GPF_CUDA(cudaFreeAsync(tmp[k], s));
            }
        GPF_CUDA(cudaStreamSynchronize(s));
    }
    return GPF_OK;
}

int launch_eos(const gpf_fv_cfg& c, const double* dens, double* out, size_t n, bool sound, cudaStream_t s) {
    EosParams e;
    int rc = eos_pack(c, &e);
    if (rc) return rc;
    if (n == 0) return GPF_OK;
    const int blocks = (int)((n + 255) / 256);
    if (sound) eos_kernel<true><<<blocks, 256, 0, s>>>(e, dens, out, n);
    else eos_kernel<false><<<blocks, 256, 0, s>>>(e, dens, out, n);
    g_launches++;
    GPF_CUDA(cudaGetLastError());
    return GPF_OK;
}

int launch_stress(const gpf_fv_cfg& c, int which, const double* q, const double* h, const double* Ls,
                  double* out, size_t n, cudaStream_t s) {
    FvParams P;
    gpf_fv_cfg c2 = c;
    if (c2.Nx < 1) c2.Nx = 1;
    if (c2.Ny < 1) c2.Ny = 1;
    int rc = pack_params(c2, &P);
    if (rc) return rc;
    if (n == 0) return GPF_OK;
    const int blocks = (int)((n + 255) / 256);
    if (Ls) stress_kernel<true><<<blocks, 256, 0, s>>>(P.visc, which, q, h, Ls, out, n);
    else stress_kernel<false><<<blocks, 256, 0, s>>>(P.visc, which, q, h, Ls, out, n);
    g_launches++;
    GPF_CUDA(cudaGetLastError());
    return GPF_OK;
}


int launch_stress_ex(const gpf_fv_cfg& c, int which, int slip_both, const double* q, const double* h, const double* Ls,
                     const double* dqx, const double* dqy, double* out, size_t n, cudaStream_t s) {
    if (which < 0 || which > 2) { set_error("which: 0 stress_bottom, 1 stress_top, 2 stress_avg"); return GPF_E_INVALID; }
    if (n == 0) return GPF_OK;
    ViscParams vp;
    vp.U = c.U; vp.V = c.V; vp.eta = c.eta; vp.zeta = c.zeta;
    vp.v1 = vp.v2 = vp.two_eta = vp.two_v2 = 0.0;
    stress_ex_kernel<<<(int)((n + 255) / 256), 256, 0, s>>>(vp, which, slip_both ? 1 : 0, q, h, Ls, dqx, dqy, out, n);
    g_launches++;
    GPF_CUDA(cudaGetLastError());
    return GPF_OK;
}

// the same routine on the host (CPU tests of the derived closed forms against the golden vectors)
int stress_ex_host(const gpf_fv_cfg& c, int which, int slip_both, const double* q, const double* h, const double* Ls,
                   const double* dqx, const double* dqy, double* out, size_t n) {
    if (which < 0 || which > 2) { set_error("which: 0 stress_bottom, 1 stress_top, 2 stress_avg"); return GPF_E_INVALID; }
    const int nc = which == 2 ? 3 : 6;
    for (size_t i = 0; i < n; ++i) {
        SlipIn in;
        in.h = h[i]; in.hx = h[n + i]; in.hy = h[2 * n + i]; in.Ls = Ls ? Ls[i] : 0.0;
        in.U = c.U; in.V = c.V; in.eta = c.eta; in.zeta = c.zeta;
        in.q0 = q[i]; in.q1 = q[n + i]; in.q2 = q[2 * n + i];
        in.dx0 = dqx ? dqx[i] : 0.0; in.dx1 = dqx ? dqx[n + i] : 0.0; in.dx2 = dqx ? dqx[2 * n + i] : 0.0;
        in.dy0 = dqy ? dqy[i] : 0.0; in.dy1 = dqy ? dqy[n + i] : 0.0; in.dy2 = dqy ? dqy[2 * n + i] : 0.0;
        double o[6];
        stress_ex_cell(which, slip_both, in, o);
        for (int k = 0; k < nc; ++k) out[k * n + i] = o[k];
    }
    return GPF_OK;
}

// ---- self-test of the branch-free reciprocal / quotient against the IEEE operations -------
__global__ void selftest_div_kernel(unsigned long long n, unsigned long long seed, unsigned long long* out) {
    unsigned long long bad_rcp = 0, bad_div = 0;
    const unsigned long long tid = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long s0 = seed ^ (tid * 0x9E3779B97F4A7C15ULL + 0x632BE59BD9B4E019ULL), s1 = seed + tid * 0xD1B54A32D192ED03ULL + 1;
    auto next = [&]() {
        unsigned long long a = s0, b = s1;
        s0 = b; a ^= a << 23; s1 = a ^ b ^ (a >> 17) ^ (b >> 26);
        return s1 + b;
    };
    auto draw = [&]() {
        unsigned long long m = next() & ((1ULL << 52) - 1);
        const unsigned mode = (unsigned)(next() % 6);
        if (mode == 1) m &= ~((1ULL << (next() % 52)) - 1);       // trailing zeros
        if (mode == 2) m |= ((1ULL << (next() % 52)) - 1);        // trailing ones
        if (mode == 3) m = (1ULL << 52) - 1 - (next() & 0xFF);    // near all ones
        if (mode == 4) m = next() & 0xFF;                         // near a power of two
        const unsigned long long e = 1023ULL + (next() % 180) - 90;   // inside the fast range
        const unsigned long long u = ((next() & 1ULL) << 63) | (e << 52) | m;
        return __longlong_as_double((long long)u);
    };
    for (unsigned long long it = tid; it < n; it += (unsigned long long)gridDim.x * blockDim.x) {
        const double a = draw(), b = draw();
        const double y = rcp_sel<true>(b);      // the reciprocal the kernels use
        if (y != __drcp_rn(b)) ++bad_rcp;
        if (div_fast(a, b, y) != __ddiv_rn(a, b)) ++bad_div;
    }
    if (bad_rcp) atomicAdd(&out[0], bad_rcp);
    if (bad_div) atomicAdd(&out[1], bad_div);
}

int launch_selftest_div(unsigned long long n, unsigned long long seed, unsigned long long* out_host, cudaStream_t s) {
    unsigned long long* d;
    GPF_CUDA(cudaMalloc(&d, 16));
    GPF_CUDA(cudaMemsetAsync(d, 0, 16, s));
    selftest_div_kernel<<<148 * 8, 256, 0, s>>>(n, seed, d);
    g_launches++;
    GPF_CUDA(cudaGetLastError());
    GPF_CUDA(cudaMemcpyAsync(out_host, d, 16, cudaMemcpyDeviceToHost, s));
    GPF_CUDA(cudaStreamSynchronize(s));
    GPF_CUDA(cudaFree(d));
    return GPF_OK;
}

}  // namespace gpf
