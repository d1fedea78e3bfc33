// GP surrogate inference on sm_100a (rows G1-G10 of SURVEY.md section 8a).
//
// Replaces, for one GP model (reference file:line in GaPFlow/):
//   models/gp.py:576-603   multi_in_single_out -> GaussianProcess(kernel, X, diag=yerr^2):
//                          K = amp*Matern32(||s.(x - x')||) + yerr^2 I,  L = chol(K)      gpf_gp_factor
//   models/gp.py:546-568   alpha = K^-1 y per output column (tinygp condition)              gpf_gp_factor
//   models/gp.py:525-535   _repredict_mean:     mean = Ks^T alpha                           gpf_gp_mean
//   models/stress.py:522-537 v_sound^2 ~ max_i d mean/d x_0 (closed form of vmap(grad))     gpf_gp_mean
//   models/gp.py:509-522   _repredict_mean_var: v = L^-1 Ks, var = amp + noise - sum v^2    gpf_gp_mean_var
//   models/gp.py:406-430   maximum_variance, argmax of the variance                         gpf_gp_mean_var
//   models/gp.py:223-232, stress.py:194-197  feature matrix (N, D)                          gpf_gp_features
//
// Design.  Training inputs are pre-scaled once (Xs = X . exp(-log_scale)); the cross-covariance
// Ks (M x N) is never written to HBM: the mean kernel evaluates it tile by tile from shared
// memory, the variance kernel evaluates the 64 x 128 block it needs directly in the accumulator
// layout of the FP64 tensor-core tiles.  L is factored by a blocked right-looking Cholesky
// (64-wide panels; panel solve and trailing update are DMMA m8n8k4 tiles), the inverses of its
// 64 x 64 diagonal blocks are kept, and every triangular solve (alpha, v) is the blocked
// substitution  x_k = inv(L_kk) (b_k - sum_{j<k} L_kj x_j)  whose inner sum is a dense FP64
// contraction on the tensor pipe.  The variance kernel is persistent: a CTA owns 128 query
// columns, walks the block rows of L in order (all CTAs stream L in the same order, so L is read
// from HBM about once per wave and served from the 126 MB L2 otherwise), keeps its v panel in a
// private HBM scratch slot and accumulates sum v^2 in registers.
// FP64 only; tcgen05 has no FP64 kind, the FP64 tensor path on sm_100a is mma.sync DMMA.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <string>
#include <vector>
#include "fv_host.h"
#include "gp_host.h"

namespace gpf {

constexpr int NB = 64;          // block size of the factorisation and of the substitutions
constexpr int MPAD = 128;       // the matrices are padded to a multiple of 128 (= the column width of the GEMM tiles): identity there
__host__ __device__ constexpr size_t pad_m(size_t m) { return (m + MPAD - 1) / MPAD * MPAD; }
constexpr int NQ = 128;         // query columns per CTA of the variance kernel
constexpr int LDT = NB + 4;     // shared-memory row stride of a 64-wide FP64 tile (2 wavefronts per fragment load)
constexpr int LDQ = NQ + 4;
constexpr double SQRT3 = 1.7320508075688772;

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Matern-3/2 on pre-scaled points
template <int D>
__device__ __forceinline__ double m32(const double* a, const double* b, double amp) {
    double d2 = 0.0;
#pragma unroll
    for (int k = 0; k < D; ++k) { const double t = a[k] - b[k]; d2 += t * t; }
    const double r = SQRT3 * sqrt(d2);
    return amp * (1.0 + r) * exp(-r);
}
__device__ __forceinline__ double m32_rt(const double* a, const double* b, int D, double amp) {
    double d2 = 0.0;
    for (int k = 0; k < D; ++k) { const double t = a[k] - b[k]; d2 += t * t; }
    const double r = SQRT3 * sqrt(d2);
    return amp * (1.0 + r) * exp(-r);
}

// ---- contracted arithmetic of the mean kernel (GPF_ARITH_FAST; north_star's GP tolerance is 1e-9 relative) --------
// exp(x), x <= 0:  x = (32 n + j) ln2/32 + f, |f| <= ln2/64;  exp(x) = 2^n T[j] (1 + f + ... + f^5/120), T[j] = 2^(j/32)
// from a 32-entry table in shared memory.  Truncation f^6/720 <= 2.3e-15; 10 FP64 instructions against ~20 of the
// correctly rounded libm routine.  The constants come from constant memory (FastK): written as literals the assembler
// re-materialises every 64-bit constant with two moves inside the loop, and this loop is bound by issue slots (r2f:
// 26 IMAD.MOV + 15 UMOV next to 48 FP64 instructions per training point).
__constant__ double GP_FASTK[10] = {
    46.16624130844683,                            // 32 / ln 2
    6755399441055744.0,                           // 1.5 * 2^52: the low word of x * rl + magic is rint(x * rl)
    -0.021660849392496573, -1.718100943346366e-15,   // -(ln 2 / 32), high and low part
    8.33333333333333322e-03, 4.16666666666666644e-02, 1.66666666666666657e-01, 0.5, 1.0, 1e-300};
struct FastK {                                    // read from constant memory: values the assembler cannot see through, so
    double rl, magic, l_hi, l_lo, c5, c4, c3, half, one, tiny;   // they are constant-bank operands or stay in registers
    __device__ __forceinline__ void init() {
        rl = GP_FASTK[0]; magic = GP_FASTK[1]; l_hi = GP_FASTK[2]; l_lo = GP_FASTK[3]; c5 = GP_FASTK[4];
        c4 = GP_FASTK[5]; c3 = GP_FASTK[6]; half = GP_FASTK[7]; one = GP_FASTK[8]; tiny = GP_FASTK[9];
    }
};
__device__ __forceinline__ double exp_neg_fast(double x, const double* T, const FastK& K) {
    // x <= 0 here; below -700 the result is 0 to every accuracy that matters: clamp on the integer pipe (for
    // negative doubles a larger magnitude is a larger high word)
    const unsigned hx = (unsigned)__double2hiint(x);
    if (hx > 0xC085E000u) x = -700.0;
    const double kd = fma(x, K.rl, K.magic);
    const int k = __double2loint(kd);
    const double kf = kd - K.magic;
    double f = fma(kf, K.l_hi, x);                // the fma forms x - k ln2/32 without intermediate rounding
    f = fma(kf, K.l_lo, f);
    double p = fma(f, K.c5, K.c4);
    p = fma(p, f, K.c3);
    p = fma(p, f, K.half);
    p = fma(p, f, K.one);
    p = fma(p, f, K.one);
    const double t = T[k & 31] * p;
    return __hiloint2double(__double2hiint(t) + ((k >> 5) << 20), __double2loint(t));
}
// sqrt(d), d > 0 (the caller starts its sum of squares at 1e-300): two corrections of d * rsqrt.approx(d) (2^-22 seed).
// One correction leaves 8e-14, which the cancellation in sum_j alpha_j k_j (|alpha| ~ 10^3) turned into 1.5e-10 on the
// mean of BASELINE config 4 (r2n); with two the contracted mean is as close to the oracle as the exact one.
__device__ __forceinline__ double sqrt_fast(double d, const FastK& K) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
    const double hy = K.half * y;
    double g = d * y;
    g = fma(fma(-g, g, d), hy, g);
    return fma(fma(-g, g, d), hy, g);
}

// warp tile 32 x 32 of a CTA tile: acc[mi][ni][2] += A(32 x 4*KK, row-major in smem) x B
//   BT = true : B given as rows n with k contiguous (C += A B^T),   Bs[n][k]
//   BT = false: B given as rows k with n contiguous (C += A B),     Bs[k][n]
//   AT = true : A given as rows k with m contiguous (C += A^T B),   As[k][m]
template <bool BT, int KK, bool AT = false>
__device__ __forceinline__ void warp_mma(double (&acc)[4][4][2], const double* As, int lda, const double* Bs, int ldb,
                                         int lane) {
    const int r = lane >> 2, c = lane & 3;
#pragma unroll
    for (int kk = 0; kk < KK; ++kk) {
        double a[4], b[4];
#pragma unroll
        for (int mi = 0; mi < 4; ++mi) a[mi] = AT ? As[(kk * 4 + c) * lda + mi * 8 + r] : As[(mi * 8 + r) * lda + kk * 4 + c];
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) b[ni] = BT ? Bs[(ni * 8 + r) * ldb + kk * 4 + c] : Bs[(kk * 4 + c) * ldb + ni * 8 + r];
#pragma unroll
        for (int mi = 0; mi < 4; ++mi)
#pragma unroll
            for (int ni = 0; ni < 4; ++ni) dmma(acc[mi][ni][0], acc[mi][ni][1], a[mi], b[ni]);
    }
}

// ------------------------------------------------------------------------------------
// K = k(X, X) + yerr^2 I on the padded Mp x Mp matrix (identity in the padding)
// ------------------------------------------------------------------------------------
__global__ void gp_build_K_kernel(const double* Xs, int M, int Mp, int D, double amp, double yerr2, double* A) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
    if (j >= Mp) return;
    double v;
    if (i >= M || j >= M) v = (i == j) ? 1.0 : 0.0;
    else {
        v = m32_rt(Xs + (size_t)i * D, Xs + (size_t)j * D, D, amp);
        if (i == j) v += yerr2;
    }
    A[(size_t)i * Mp + j] = v;
}

// ------------------------------------------------------------------------------------
// Cholesky step k, part 1: factor the diagonal block in shared memory, write L_kk back
// and inv(L_kk) to Dinv[k].  One CTA of 256 threads.  info != 0: not positive definite.
// ------------------------------------------------------------------------------------
constexpr size_t POTRF_SMEM = (size_t)(2 * NB * (NB + 1) + 4 * NB) * 8;     // L_kk, its inverse, two pivot columns, two pivot rows

// 1 / sqrt(d) to full precision: hardware seed (2^-22) and two Newton steps
__device__ __forceinline__ double rsqrt_full(double d) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
    double e = fma(-d * y, y, 1.0);
    y = fma(y * e, fma(e, 0.375, 0.5), y);         // y (1 + e/2 + 3 e^2/8)
    e = fma(-d * y, y, 1.0);
    return fma(y * e, 0.5, y);
}

// Columns 16 Q1 .. 16 Q1 + 15 of the in-register factor-and-invert sweep (see gp_potrf_diag_kernel).  Q1 is a
// template parameter so that every register index is static and the cells that cannot change any more (block rows
// above Q1; block columns of a left of Q1, of b right of Q1) drop out of the code: only the 16-wide band that
// contains the pivot needs a run-time comparison.
template <int Q1>
__device__ __forceinline__ void potrf_phase(double (&a)[4][4], double (&b)[4][4], double (*Lm)[NB + 1], double (*X)[NB + 1],
                                            double* colbuf, double* rowbuf, int ty, int tx, int k, int* info) {
    for (int jt = 0; jt < 16; ++jt) {
        const int j = 16 * Q1 + jt;
        const double* cb = colbuf + (j & 1) * NB;
        const double* rb = rowbuf + (j & 1) * NB;
        const double d = cb[j];
        if (ty == 0 && tx == 0 && !(d > 0.0)) atomicExch(info, k * NB + j + 1);
        const double inv = rsqrt_full(d);
        double lr[4], lc[4], xr[4];
#pragma unroll
        for (int p = Q1; p < 4; ++p) { lr[p] = cb[ty + 16 * p] * inv; lc[p] = cb[tx + 16 * p] * inv; }
#pragma unroll
        for (int q = 0; q <= Q1; ++q) xr[q] = rb[tx + 16 * q] * inv;
        if (tx == jt) {                             // owners of column j: the finished column of L (zeros above are preset)
            if (ty >= jt) Lm[ty + 16 * Q1][j] = ty > jt ? lr[Q1] : d * inv;
#pragma unroll
            for (int p = Q1 + 1; p < 4; ++p) Lm[ty + 16 * p][j] = lr[p];
        }
        if (ty == jt) {                             // owners of row j: the finished row of the inverse
#pragma unroll
            for (int q = 0; q <= Q1; ++q) X[j][tx + 16 * q] = xr[q];
        }
        // rows below the pivot: a -= l_i l_c right of the pivot column, b -= l_i x_c (x_c = 0 right of the pivot)
        const bool row_band = ty > jt, col_band = tx > jt;
#pragma unroll
        for (int p = Q1; p < 4; ++p) {
            if (p == Q1 && !row_band) continue;
            if (col_band) a[p][Q1] = fma(-lr[p], lc[Q1], a[p][Q1]);
#pragma unroll
            for (int q = Q1 + 1; q < 4; ++q) a[p][q] = fma(-lr[p], lc[q], a[p][q]);
#pragma unroll
            for (int q = 0; q <= Q1; ++q) b[p][q] = fma(-lr[p], xr[q], b[p][q]);
        }
        // publish the next pivot column of a and pivot row of b
        double* nc = colbuf + ((j + 1) & 1) * NB;
        double* nr = rowbuf + ((j + 1) & 1) * NB;
        if (jt < 15) {
            if (tx == jt + 1) {
#pragma unroll
                for (int p = Q1; p < 4; ++p) nc[ty + 16 * p] = a[p][Q1];
            }
            if (ty == jt + 1) {
#pragma unroll
                for (int q = 0; q <= Q1; ++q) nr[tx + 16 * q] = b[Q1][q];
            }
        } else if (Q1 < 3) {
            constexpr int QN = Q1 < 3 ? Q1 + 1 : 3;
            if (tx == 0) {
#pragma unroll
                for (int p = QN; p < 4; ++p) nc[ty + 16 * p] = a[p][QN];
            }
            if (ty == 0) {
#pragma unroll
                for (int q = 0; q <= QN; ++q) nr[tx + 16 * q] = b[QN][q];
            }
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) gp_potrf_diag_kernel(double* A, int Mp, int k, double* Dinv, int* info) {
    extern __shared__ double potrf_sm[];
    double (*X)[NB + 1] = reinterpret_cast<double (*)[NB + 1]>(potrf_sm);                   // inv(L_kk)
    double (*Lm)[NB + 1] = reinterpret_cast<double (*)[NB + 1]>(potrf_sm + NB * (NB + 1));  // L_kk
    double* colbuf = potrf_sm + 2 * NB * (NB + 1);                                          // [2][NB] the pivot column of A, raw
    double* rowbuf = colbuf + 2 * NB;                                                       // [2][NB] the pivot row of B, raw
    const int tid = threadIdx.x;
    double* Akk = A + (size_t)k * NB * Mp + (size_t)k * NB;
    // Factor and invert in one sweep, both blocks in registers: thread (ty, tx) of 16 x 16 owns the cells
    // (ty + 16 p, tx + 16 q) of the working copy a of A_kk and of b, which starts as the identity and carries the
    // forward substitution L X = I along (row j of X is row j of b over L_jj once the columns before j are eliminated).
    // Per column ONE barrier: the owners of the next pivot column of a and pivot row of b publish them (raw) in shared
    // memory, every thread forms 1 / sqrt(d_jj) itself and updates its cells with independent multiply-adds.
    // (With the block in shared memory, three barriers per column and a separate inverse: 69 us per block, 11 ms per
    // factorisation at M = 10^4 -- the serial chain that the trailing updates cannot hide.)
    const int ty = tid >> 4, tx = tid & 15;
    double a[4][4], b[4][4];
#pragma unroll
    for (int p = 0; p < 4; ++p)
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            a[p][q] = Akk[(size_t)(ty + 16 * p) * Mp + tx + 16 * q];
            b[p][q] = (ty + 16 * p == tx + 16 * q) ? 1.0 : 0.0;
        }
    for (int e = tid; e < 2 * NB * (NB + 1); e += 256) potrf_sm[e] = 0.0;     // X and Lm: the zero triangles
    if (tx == 0) {
#pragma unroll
        for (int p = 0; p < 4; ++p) colbuf[ty + 16 * p] = a[p][0];
    }
    if (ty == 0) rowbuf[tx] = b[0][0];
    __syncthreads();
    potrf_phase<0>(a, b, Lm, X, colbuf, rowbuf, ty, tx, k, info);
    potrf_phase<1>(a, b, Lm, X, colbuf, rowbuf, ty, tx, k, info);
    potrf_phase<2>(a, b, Lm, X, colbuf, rowbuf, ty, tx, k, info);
    potrf_phase<3>(a, b, Lm, X, colbuf, rowbuf, ty, tx, k, info);
    double* Dk = Dinv + (size_t)k * NB * NB;
    for (int e = tid; e < NB * NB; e += 256) {
        const int i = e / NB, c = e % NB;
        Akk[(size_t)i * Mp + c] = Lm[i][c];
        Dk[e] = X[i][c];
    }
}

// ------------------------------------------------------------------------------------
// Cholesky step k, part 2: panel solve  A[i,k] <- A[i,k] inv(L_kk)^T  for block rows i > k.
// One CTA (128 threads, 2 x 2 warps of 32 x 32) per block row.
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) gp_panel_kernel(double* A, int Mp, int k, const double* Dinv) {
    extern __shared__ double sm[];
    double* As = sm;
    double* Bs = sm + NB * LDT;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5, wm = w >> 1, wn = w & 1;
    const int i = k + 1 + blockIdx.x;
    double* Aik = A + (size_t)i * NB * Mp + (size_t)k * NB;
    const double* Dk = Dinv + (size_t)k * NB * NB;
    for (int e = tid; e < NB * NB; e += 128) {
        As[(e / NB) * LDT + e % NB] = Aik[(size_t)(e / NB) * Mp + e % NB];
        Bs[(e / NB) * LDT + e % NB] = Dk[e];
    }
    __syncthreads();
    double acc[4][4][2] = {};
    warp_mma<true, NB / 4>(acc, As + wm * 32 * LDT, LDT, Bs + wn * 32 * LDT, LDT, lane);
    const int r = lane >> 2, c = lane & 3;
#pragma unroll
    for (int mi = 0; mi < 4; ++mi)
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) {
            double* p = Aik + (size_t)(wm * 32 + mi * 8 + r) * Mp + wn * 32 + ni * 8 + c * 2;
            p[0] = acc[mi][ni][0];
            p[1] = acc[mi][ni][1];
        }
}

// ------------------------------------------------------------------------------------
// Cholesky step k, part 3: update of the rest of the OUTER panel  A[i,j] -= A[i,k] A[j,k]^T  for the block columns
// k < j < jend of the 256-wide outer panel and block rows i >= j (2D grid: i = k + 1 + x, j = k + 1 + y).  Everything
// right of the outer panel is updated once per outer panel by the GEMM engine (GEMM_SYRK, k-depth 256).
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) gp_syrk_kernel(double* A, int Mp, int k) {
    extern __shared__ double sm[];
    double* As = sm;
    double* Bs = sm + NB * LDT;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5, wm = w >> 1, wn = w & 1;
    const int i = k + 1 + blockIdx.x, j = k + 1 + blockIdx.y;
    if (j > i) return;
    const double* Aik = A + (size_t)i * NB * Mp + (size_t)k * NB;
    const double* Ajk = A + (size_t)j * NB * Mp + (size_t)k * NB;
    for (int e = tid; e < NB * NB; e += 128) {
        As[(e / NB) * LDT + e % NB] = Aik[(size_t)(e / NB) * Mp + e % NB];
        Bs[(e / NB) * LDT + e % NB] = Ajk[(size_t)(e / NB) * Mp + e % NB];
    }
    __syncthreads();
    double acc[4][4][2] = {};
    warp_mma<true, NB / 4>(acc, As + wm * 32 * LDT, LDT, Bs + wn * 32 * LDT, LDT, lane);
    double* Cij = A + (size_t)i * NB * Mp + (size_t)j * NB;
    const int r = lane >> 2, c = lane & 3;
#pragma unroll
    for (int mi = 0; mi < 4; ++mi)
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) {
            double* q = Cij + (size_t)(wm * 32 + mi * 8 + r) * Mp + wn * 32 + ni * 8 + c * 2;
            q[0] -= acc[mi][ni][0];
            q[1] -= acc[mi][ni][1];
        }
}

// ------------------------------------------------------------------------------------
// alpha = K^-1 Y by blocked, right-looking substitution: step k solves block k with the stored
// inverse of L_kk and at once removes its contribution from every remaining block, one CTA per
// remaining block (the left-looking form -- one CTA streaming a whole block row of L per step --
// took 174 of the 207 ms of a factorisation at M = 10^4).
//   forward  (k = 0 .. nb-1):  z_k = Dinv_k w_k;    w_i -= L_ik z_k     for i > k
//   backward (k = nb-1 .. 0):  a_k = Dinv_k^T z_k;  z_j -= L_kj^T a_k   for j < k
// Every CTA of a step recomputes the 64 x 64 solve (32 KB of Dinv from L2); the vectors are
// (nrhs, Mp) row-major, nrhs <= 2.  W is destroyed, Z holds z and is destroyed by the backward sweep.
// ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) gp_trsv_fwd_kernel(const double* L, int Mp, int k, const double* Dinv, double* W,
                                                          double* Z, int nrhs) {
    __shared__ double r[2][NB], z[2][NB];
    const int tid = threadIdx.x;
    const double* Dk = Dinv + (size_t)k * NB * NB;
    if (tid < NB * nrhs) r[tid / NB][tid % NB] = W[(size_t)(tid / NB) * Mp + k * NB + tid % NB];
    __syncthreads();
    if (tid < NB * nrhs) {
        const int o = tid / NB, row = tid % NB;
        double s = 0.0;
        for (int c = 0; c <= row; ++c) s += Dk[row * NB + c] * r[o][c];
        z[o][row] = s;
        if (blockIdx.x == 0) Z[(size_t)o * Mp + k * NB + row] = s;
    }
    if (blockIdx.x == 0) return;
    __syncthreads();
    const int row = (k + blockIdx.x) * NB + (tid >> 2), part = tid & 3;
    const double* Lr = L + (size_t)row * Mp + k * NB + part * 16;
    for (int o = 0; o < nrhs; ++o) {
        double s = 0.0;
#pragma unroll
        for (int c = 0; c < 16; ++c) s += Lr[c] * z[o][part * 16 + c];
        s += __shfl_xor_sync(0xffffffffu, s, 1);
        s += __shfl_xor_sync(0xffffffffu, s, 2);
        if (part == 0) W[(size_t)o * Mp + row] -= s;
    }
}
__global__ void __launch_bounds__(256) gp_trsv_bwd_kernel(const double* L, int Mp, int k, const double* Dinv, double* Z,
                                                          double* A, int nrhs) {
    __shared__ double r[2][NB], a[2][NB];
    __shared__ double part[4][2][NB];
    const int tid = threadIdx.x;
    const double* Dk = Dinv + (size_t)k * NB * NB;
    if (tid < NB * nrhs) r[tid / NB][tid % NB] = Z[(size_t)(tid / NB) * Mp + k * NB + tid % NB];
    __syncthreads();
    if (tid < NB * nrhs) {
        const int o = tid / NB, col = tid % NB;
        double s = 0.0;
        for (int row = col; row < NB; ++row) s += Dk[row * NB + col] * r[o][row];     // Dinv^T
        a[o][col] = s;
        if ((int)blockIdx.x == k) A[(size_t)o * Mp + k * NB + col] = s;
    }
    if ((int)blockIdx.x == k) return;
    __syncthreads();
    const int j = blockIdx.x, c = tid & 63, g = tid >> 6;
    const double* Lb = L + (size_t)(k * NB + g * 16) * Mp + j * NB + c;
    for (int o = 0; o < nrhs; ++o) {
        double s = 0.0;
#pragma unroll
        for (int rr = 0; rr < 16; ++rr) s += Lb[(size_t)rr * Mp] * a[o][g * 16 + rr];
        part[g][o][c] = s;
    }
    __syncthreads();
    if (tid < NB * nrhs) {
        const int o = tid / NB, cc = tid % NB;
        Z[(size_t)o * Mp + j * NB + cc] -= part[0][o][cc] + part[1][o][cc] + part[2][o][cc] + part[3][o][cc];
    }
}

// ------------------------------------------------------------------------------------
// mean (n_out, N) = alpha^T Ks, and d mean_0 / d x_0 (max over the queries of this block).
// One thread per query, training points staged through shared memory.  The loop is FP64-issue bound
// (sqrt ~10 and exp ~20 instructions of ~50 per kernel evaluation), so the variants drop what a call does
// not need: NOUT = 1 the second output, GRAD = false the slope sum, D = 1 the square root
// (sqrt(RN(t*t)) == |t| exactly unless t*t leaves the normal range, where both give 0 or inf alike).
// ------------------------------------------------------------------------------------
// Two queries per thread (Q = 2: cells i and i + 256 of the block's 512): every shared-memory load of a
// training point and of its alpha serves two kernel evaluations, which gives the ~13 % of issue slots
// the loads took (profiles/r1C_ncu_coupled_summary.csv) back to the FP64 pipe.
template <int D, int NOUT, bool GRAD, bool FASTM>
__global__ void __launch_bounds__(256) gp_mean_kernel(const double* Xs, const double* alpha, int M, int Mp, int nout,
                                                      GpKern kp, const double* Xtest, long long N, long long off,
                                                      long long cnt, double* mean, double* gmax_part) {
    constexpr int TM = 256, Q = 2;
    __shared__ double sx[TM * D];
    __shared__ double sa[2][TM];
    __shared__ double sred[8];
    __shared__ double sT[32];
    // contracted variant: inputs carry the factor sqrt(3), so r = sqrt(sum t^2) needs no multiplication
    constexpr double PRE = FASTM ? SQRT3 : 1.0;
    if (FASTM && threadIdx.x < 32) sT[threadIdx.x] = exp2((double)threadIdx.x / 32.0);
    FastK K;
    if (FASTM) K.init();
    long long i[Q];
    bool valid[Q];
    double x[Q][D];
    double m0[Q], m1[Q], g0[Q];
#pragma unroll
    for (int q = 0; q < Q; ++q) {
        i[q] = off + (long long)blockIdx.x * (256 * Q) + q * 256 + threadIdx.x;
        valid[q] = i[q] < off + cnt;
#pragma unroll
        for (int k = 0; k < D; ++k) x[q][k] = valid[q] ? Xtest[i[q] * D + k] * kp.s[k] * PRE : 0.0;
        m0[q] = m1[q] = g0[q] = 0.0;
    }
    for (int t0 = 0; t0 < M; t0 += TM) {
        const int n = min(TM, M - t0);
        __syncthreads();
        for (int e = threadIdx.x; e < n * D; e += 256) sx[e] = Xs[(size_t)t0 * D + e] * PRE;
        for (int e = threadIdx.x; e < n; e += 256) {
            sa[0][e] = alpha[t0 + e];
            if (NOUT > 1) sa[1][e] = alpha[(size_t)Mp + t0 + e];
        }
        __syncthreads();
        for (int j = 0; j < n; ++j) {
            double xt[D];
#pragma unroll
            for (int k = 0; k < D; ++k) xt[k] = sx[j * D + k];
            const double a0 = sa[0][j];
            const double a1 = NOUT > 1 ? sa[1][j] : 0.0;
#pragma unroll
            for (int q = 0; q < Q; ++q) {
                double r;
                if constexpr (D == 1) {
                    r = PRE == 1.0 ? SQRT3 * fabs(x[q][0] - xt[0]) : fabs(x[q][0] - xt[0]);
                } else {
                    double d2 = FASTM ? K.tiny : 0.0;
#pragma unroll
                    for (int k = 0; k < D; ++k) { const double t = x[q][k] - xt[k]; d2 += t * t; }
                    r = FASTM ? sqrt_fast(d2, K) : SQRT3 * sqrt(d2);
                }
                const double e = FASTM ? exp_neg_fast(-r, sT, K) : exp(-r);
                const double kv = FASTM ? fma(r, e, e) : (1.0 + r) * e;
                m0[q] += a0 * kv;
                if (NOUT > 1) m1[q] += a1 * kv;
                if (GRAD) g0[q] += a0 * ((x[q][0] - xt[0]) * e);
            }
        }
    }
    double g = -INFINITY;
#pragma unroll
    for (int q = 0; q < Q; ++q)
        if (valid[q]) {
            mean[i[q]] = kp.amp * m0[q];
            if (NOUT > 1) mean[N + i[q]] = kp.amp * m1[q];
            // d/dx_0 in the (unscaled) normalised input; the contracted variant's differences carry sqrt(3)
            g = fmax(g, (FASTM ? -SQRT3 : -3.0) * kp.amp * kp.s[0] * g0[q]);
        }
    if (!GRAD) return;
    // block max (NaN-propagating like jnp.max is not needed: a NaN gradient makes the mean NaN too)
#pragma unroll
    for (int oo = 16; oo > 0; oo >>= 1) g = fmax(g, __shfl_xor_sync(0xffffffffu, g, oo));
    if ((threadIdx.x & 31) == 0) sred[threadIdx.x >> 5] = g;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int ww = 1; ww < 8; ++ww) g = fmax(g, sred[ww]);
        gmax_part[blockIdx.x] = g;
    }
}

__global__ void gp_max_reduce_kernel(const double* part, int n, double* out) {
    __shared__ double s[256];
    double g = -INFINITY;
    for (int e = threadIdx.x; e < n; e += 256) g = fmax(g, part[e]);
    s[threadIdx.x] = g;
    __syncthreads();
    for (int st = 128; st > 0; st >>= 1) {
        if (threadIdx.x < st) s[threadIdx.x] = fmax(s[threadIdx.x], s[threadIdx.x + st]);
        __syncthreads();
    }
    if (threadIdx.x == 0) *out = s[0];
}

// ------------------------------------------------------------------------------------
// var (N) = amp + noise - sum_m v[m]^2,  v = L^-1 Ks, by blocked forward substitution.
// Persistent CTAs (256 threads = 2 x 4 warps of 32 x 32): a CTA owns NQ = 128 query columns.
// ------------------------------------------------------------------------------------
constexpr int VBK = 16;                       // k-depth of one pipeline stage
constexpr int VST = 3;                        // cp.async stages
constexpr int LDA_V = VBK + 4;                // row stride of the staged L sub-tile: = 4 mod 16 doubles, so the 16 lanes of a
                                              // half-warp (4 rows x 4 k) hit 16 different 8-byte banks (r2a: VBK + 2 gave 2-way
                                              // conflicts on every A-fragment load, 29 % of the kernel's shared-memory wavefronts)
constexpr size_t VAR_STAGE = (size_t)(NB * LDA_V + VBK * LDQ) * 8;
constexpr size_t VAR_SOLVE = (size_t)(NB * LDQ) * 8;                          // R = Ks - acc, the B operand of the block solve
constexpr size_t VAR_MAIN = (VST * VAR_STAGE > VAR_SOLVE) ? VST * VAR_STAGE : VAR_SOLVE;
constexpr int GP_MAXD = 8;
constexpr size_t VAR_SMEM = VAR_MAIN + (size_t)(NQ * GP_MAXD + NB * GP_MAXD + NQ * 2) * 8;
constexpr int VAR_CTAS_PER_SM = 2;            // 95 KB and <= 128 registers per CTA: two CTAs share an SM, so the tensor pipe has
                                              // work while the other CTA sits at a barrier or in its block solve (r2a, one CTA per
                                              // SM: DMMA pipe 72 % busy, stalls = barrier + cp.async wait)

// A-fragment loads of the 64 x 64 block solve straight from global memory (inv(L_kk), 32 KB, L2-resident: every CTA
// is at a similar block row); keeps the solve's A operand out of shared memory
__device__ __forceinline__ void warp_mma_gA(double (&acc)[4][4][2], const double* Ag, int lda, const double* Bs, int ldb, int lane) {
    const int r = lane >> 2, c = lane & 3;
#pragma unroll 4
    for (int kk = 0; kk < NB / 4; ++kk) {
        double a[4], b[4];
#pragma unroll
        for (int mi = 0; mi < 4; ++mi) a[mi] = __ldg(Ag + (size_t)(mi * 8 + r) * lda + kk * 4 + c);
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) b[ni] = Bs[(kk * 4 + c) * ldb + ni * 8 + r];
#pragma unroll
        for (int mi = 0; mi < 4; ++mi)
#pragma unroll
            for (int ni = 0; ni < 4; ++ni) dmma(acc[mi][ni][0], acc[mi][ni][1], a[mi], b[ni]);
    }
}

__global__ void __launch_bounds__(256, VAR_CTAS_PER_SM) gp_var_kernel(const double* L, const double* Dinv, const double* Xs, int M, int Mp,
                                                        GpKern kp, const double* Xtest, long long N, const double* noise,
                                                        double noise_scalar, double* var, double* scratch,
                                                        double* vmax_part, long long* imax_part) {
    extern __shared__ __align__(16) unsigned char smraw[];
    double* sm = reinterpret_cast<double*>(smraw);
    double* sxq = reinterpret_cast<double*>(smraw + VAR_MAIN);     // scaled query features [NQ][D]
    double* sxt = sxq + NQ * GP_MAXD;                               // scaled training features of block kb [NB][D]
    double* scol = sxt + NB * GP_MAXD;                              // column sums from the two warp rows [2][NQ]
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5, wm = w >> 2, wn = w & 3;
    const int r = lane >> 2, c = lane & 3;
    const int D = kp.D;
    const int nb = Mp / NB;
    double* V = scratch + (size_t)blockIdx.x * Mp * NQ;             // this CTA's v panel (Mp x NQ)
    const long long ntiles = (N + NQ - 1) / NQ;
    double best = -INFINITY;
    long long besti = 0;

    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long q0 = tile * NQ;
        __syncthreads();
        for (int e = tid; e < NQ * D; e += 256) {
            const long long qi = q0 + e / D;
            sxq[(e / D) * GP_MAXD + e % D] = (qi < N) ? Xtest[qi * D + e % D] * kp.s[e % D] : 0.0;
        }
        double colsum[4][2] = {};
        for (int kb = 0; kb < nb; ++kb) {
            // ---- acc = sum_{k < kb*NB} L[kb rows][k] * V[k][cols], 3-stage cp.async pipeline ----
            double acc[4][4][2] = {};
            const int nk = kb * NB / VBK;
            const double* Lrow = L + (size_t)kb * NB * Mp;
            // L sub-tile: 64 rows x 16 doubles = 512 16-byte chunks (2 per thread); V sub-tile: 16 rows x 128 = 1024 chunks (4)
            const int lrow = tid >> 3, lch = tid & 7, vrow = tid >> 6, vch = tid & 63;
            auto stage_load = [&](int s, int kt) {
                double* As = sm + (size_t)s * (VAR_STAGE / 8);
                double* Bs = As + NB * LDA_V;
                const double* lg = Lrow + (size_t)kt * VBK + lch * 2;
#pragma unroll
                for (int h = 0; h < 2; ++h) cp_async16(As + (lrow + 32 * h) * LDA_V + lch * 2, lg + (size_t)(lrow + 32 * h) * Mp);
                const double* vg = V + (size_t)(kt * VBK) * NQ + vch * 2;
#pragma unroll
                for (int h = 0; h < 4; ++h) cp_async16(Bs + (vrow + 4 * h) * LDQ + vch * 2, vg + (size_t)(vrow + 4 * h) * NQ);
            };
            __syncthreads();                      // previous block's solve buffer / V stores are done
            for (int s = 0; s < VST - 1; ++s) {
                if (s < nk) stage_load(s, s);
                cp_async_commit();
            }
            for (int kt = 0; kt < nk; ++kt) {
                cp_async_wait<VST - 2>();
                __syncthreads();
                const int nx = kt + VST - 1;
                if (nx < nk) stage_load(nx % VST, nx);
                cp_async_commit();
                const double* As = sm + (size_t)(kt % VST) * (VAR_STAGE / 8);
                const double* Bs = As + NB * LDA_V;
                warp_mma<false, VBK / 4>(acc, As + wm * 32 * LDA_V, LDA_V, Bs + wn * 32, LDQ, lane);
            }
            cp_async_wait<0>();
            __syncthreads();
            // ---- R = Ks[kb] - acc in the accumulator layout, staged as the B operand of the block solve.  The sums go
            //      first so that the accumulators are dead while the kernel function is evaluated ----
            for (int e = tid; e < NB * D; e += 256) {
                const int m = kb * NB + e / D;
                sxt[(e / D) * GP_MAXD + e % D] = (m < M) ? Xs[(size_t)m * D + e % D] : 0.0;
            }
            double* Rs = sm;                              // [NB][LDQ]
#pragma unroll
            for (int mi = 0; mi < 4; ++mi)
#pragma unroll
                for (int ni = 0; ni < 4; ++ni) {
                    const int m = wm * 32 + mi * 8 + r, n = wn * 32 + ni * 8 + c * 2;
                    *reinterpret_cast<double2*>(Rs + m * LDQ + n) = make_double2(-acc[mi][ni][0], -acc[mi][ni][1]);
                }
            __syncthreads();                              // sxt complete (the Rs elements are this thread's own)
#pragma unroll 1
            for (int mi = 0; mi < 4; ++mi) {
                const int m = wm * 32 + mi * 8 + r;
                if (kb * NB + m >= M) continue;           // padding rows: R = -acc
#pragma unroll 2
                for (int ni = 0; ni < 4; ++ni)
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const int n = wn * 32 + ni * 8 + c * 2 + h;
                        Rs[m * LDQ + n] += m32_rt(sxt + m * GP_MAXD, sxq + n * GP_MAXD, D, kp.amp);
                    }
            }
            __syncthreads();
            // ---- V[kb] = inv(L_kk) R ----
            double vv[4][4][2] = {};
            warp_mma_gA(vv, Dinv + (size_t)kb * NB * NB + (size_t)wm * 32 * NB, NB, Rs + wn * 32, LDQ, lane);
#pragma unroll
            for (int mi = 0; mi < 4; ++mi)
#pragma unroll
                for (int ni = 0; ni < 4; ++ni) {
                    const int m = kb * NB + wm * 32 + mi * 8 + r, n = wn * 32 + ni * 8 + c * 2;
                    *reinterpret_cast<double2*>(V + (size_t)m * NQ + n) = make_double2(vv[mi][ni][0], vv[mi][ni][1]);
                    colsum[ni][0] += vv[mi][ni][0] * vv[mi][ni][0];
                    colsum[ni][1] += vv[mi][ni][1] * vv[mi][ni][1];
                }
        }
        // ---- fold the column sums: 8 lanes share a column (xor 4, 8, 16), 2 warp rows through smem ----
#pragma unroll
        for (int ni = 0; ni < 4; ++ni)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                double s = colsum[ni][h];
                s += __shfl_xor_sync(0xffffffffu, s, 4);
                s += __shfl_xor_sync(0xffffffffu, s, 8);
                s += __shfl_xor_sync(0xffffffffu, s, 16);
                if (r == 0) scol[wm * NQ + wn * 32 + ni * 8 + c * 2 + h] = s;
            }
        __syncthreads();
        if (tid < NQ) {
            const long long qi = q0 + tid;
            if (qi < N) {
                const double nz = noise ? noise[qi] : noise_scalar;
                const double v = kp.amp + nz - (scol[tid] + scol[NQ + tid]);
                var[qi] = v;
                if (v > best || v != v) { if (!(best != best)) { best = v; besti = qi; } }   // first maximum; NaN sticks
            }
        }
    }
    // per-CTA (max, first index); tid < NQ hold candidates
    __shared__ double sb[NQ];
    __shared__ long long si[NQ];
    if (tid < NQ) { sb[tid] = best; si[tid] = besti; }
    __syncthreads();
    if (tid == 0) {
        double b = -INFINITY; long long bi = 0x7fffffffffffffffLL; bool nanv = false;
        for (int t = 0; t < NQ; ++t) {
            const double v = sb[t];
            if (v != v) { if (!nanv || si[t] < bi) { nanv = true; b = v; bi = si[t]; } continue; }
            if (nanv) continue;
            if (v > b || (v == b && si[t] < bi)) { b = v; bi = si[t]; }
        }
        vmax_part[blockIdx.x] = b;
        imax_part[blockIdx.x] = bi;
    }
}

__global__ void gp_argmax_reduce_kernel(const double* vpart, const long long* ipart, int n, double* vout, long long* iout) {
    if (threadIdx.x != 0) return;
    double b = -INFINITY; long long bi = 0x7fffffffffffffffLL; bool nanv = false;
    for (int t = 0; t < n; ++t) {
        const double v = vpart[t];
        if (v != v) { if (!nanv || ipart[t] < bi) { nanv = true; b = v; bi = ipart[t]; } continue; }
        if (nanv) continue;
        if (v > b || (v == b && ipart[t] < bi)) { b = v; bi = ipart[t]; }
    }
    *vout = b;
    *iout = bi;
}

// ------------------------------------------------------------------------------------
// Feature matrix (gp.py:223-232, stress.py:194-197): row i*(Ny+2)+j, columns = active dims of
// [rho, jx, jy, h, dh/dx, dh/dy, extra] / X_scale, read from the FV solver's own storage.
// ------------------------------------------------------------------------------------
struct FeatArgs { int D; int dims[GP_MAXD]; double inv_dummy; double scale[GP_MAXD]; };
__global__ void gp_features_kernel(const double* q, const double* topo, const double* extra, size_t plane, int pitch,
                                   int Nx2, int Ny2, FeatArgs fa, double* X) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)Nx2 * Ny2) return;
    const int i = (int)(idx / Ny2), j = (int)(idx % Ny2);
    const size_t a = (size_t)i * pitch + (j + 1);
    for (int k = 0; k < fa.D; ++k) {
        const int f = fa.dims[k];
        double v;
        if (f < 3) v = q[f * plane + a];
        else if (f < 6) v = topo[(f - 3) * plane + a];
        else v = extra[a];
        X[idx * fa.D + k] = v / fa.scale[k];
    }
}

// =====================================================================================
// Hyperparameter training (row G9, gp.py:290-335): negative log marginal likelihood and its
// analytic gradient at the parameters of the last factorisation.
//   loss = sum_o [ 1/2 y_o^T alpha_o ] + n_out (sum_i log L_ii + M/2 log 2 pi)
//   d loss / d theta = 1/2 sum_ij W_ij dK_ij/dtheta,  W = n_out K^-1 - sum_o alpha_o alpha_o^T
//   dk/dlog_amp = k,   dk/dlog_scale_d = 3 amp exp(-sqrt3 r) (s_d (x_d - x'_d))^2
// K^-1 = Linv^T Linv with Linv = L^-1 built block column by block column from the stored
// inverses of the diagonal blocks.
// =====================================================================================
__device__ __forceinline__ void load_tile(double* dst, const double* src, int ld, int tid, int nthr) {
    for (int e = tid; e < NB * NB; e += nthr) dst[(e / NB) * LDT + e % NB] = src[(size_t)(e / NB) * ld + e % NB];
}
__device__ __forceinline__ void load_tile_t(double* dst, const double* src, int ld, int tid, int nthr) {
    for (int e = tid; e < NB * NB; e += nthr) dst[(e % NB) * LDT + e / NB] = src[(size_t)(e / NB) * ld + e % NB];
}

// ------------------------------------------------------------------------------------
// One DMMA GEMM engine for the dense level-3 work of training: a CTA (256 threads, 2 x 4 warps of 32 x 32) forms a
// 64 x 128 tile  sum_{k0 <= k < k1} A(m, k) B(k, n)  through a 3-stage cp.async pipeline of 16-deep slices -- the
// main loop of the variance kernel with the operand layouts as template parameters:
//   AT: A stored as [k][m] (the product uses A^T),  BT: B stored as [n][k] (the product uses B^T).
// Two CTAs share an SM.  What the tile is, which k range is non-zero and what happens to the result depends on
// the job (tile coordinates from blockIdx, no tile list in memory):
//   GEMM_TRTRI_T   T = L21 X11        one level of the recursive inverse, first product   (NN)  -> T
//   GEMM_TRTRI_X   X21 = -X22 T       second product                                      (NN)  -> X
//   GEMM_LAUUM     Kinv = X^T X       lower triangle                                      (TN)  -> Kinv
//   GEMM_SYRK      A22 -= A21 A21^T   trailing update of the blocked Cholesky, k-depth = the 256-wide outer panel (NT)
// ------------------------------------------------------------------------------------
enum { GEMM_TRTRI_T = 0, GEMM_TRTRI_X = 1, GEMM_LAUUM = 2, GEMM_SYRK = 3 };
constexpr int GK = 16;                                   // k-depth of a pipeline slice
constexpr int GST = 3;
constexpr int G_LDA = GK + 4, G_LDAT = NB + 4, G_LDB = NQ + 4, G_LDBT = GK + 4;
constexpr size_t GEMM_STAGE = (size_t)(NB * G_LDA + NQ * G_LDBT) * 8;      // the larger of the layouts: 30 720 bytes
constexpr size_t GEMM_SMEM = GST * GEMM_STAGE;
static_assert((size_t)(GK * G_LDAT + GK * G_LDB) * 8 <= GEMM_STAGE, "stage size");

struct GemmJob {
    int mode;
    int Mp;              // leading dimension of every matrix
    int h;               // TRTRI: blocks (of 64) in the left part of a group at this level
    const double* L;     // TRTRI_T: the Cholesky factor
    double* X;           // the inverse under construction (TRTRI), its input (LAUUM)
    double* T;           // TRTRI: scratch with the layout of X;  LAUUM: Kinv
    int kb0, kb1;        // SYRK: the outer panel's block columns [kb0, kb1) (the k range); rows and columns from block kb1
    int ct0;             // SYRK: first 128-column tile (relative to block kb1) of this launch
};

template <bool AT, bool BT>
__global__ void __launch_bounds__(256, 2) gp_gemm_kernel(const GemmJob J) {
    extern __shared__ __align__(16) unsigned char gsm[];
    double* sm = reinterpret_cast<double*>(gsm);
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5, wm = w >> 2, wn = w & 3;
    const int r = lane >> 2, c = lane & 3;
    const int ld = J.Mp, nb = J.Mp / NB;
    // ---- the tile: rows m0 .. m0+63, columns n0 .. n0+127, k in [k0, k1) of A (lda) and B (ldb) -> C ----
    int m0, n0, k0, k1;
    const double *A, *B;
    double* C;
    bool negate = false, subtract = false;
    if (J.mode == GEMM_SYRK) {
        // A[m, n] -= sum_{k in panel} A[m, k] A[n, k]: lower triangle of the trailing matrix (tiles that straddle the
        // diagonal also touch entries above it, which nothing reads)
        m0 = (J.kb1 + blockIdx.x) * NB; n0 = J.kb1 * NB + (J.ct0 + blockIdx.y) * NQ;
        if (n0 > m0 + NB - 1 || n0 >= J.Mp) return;
        k0 = J.kb0 * NB; k1 = J.kb1 * NB;
        A = J.X; B = J.X; C = J.X;
        subtract = true;
    } else if (J.mode == GEMM_LAUUM) {
        // Kinv[m, n] = sum_{k >= m} X[k, m] X[k, n] for n0 <= m0 + 63 (X is lower triangular: X[k, .] = 0 left of k)
        m0 = blockIdx.x * NB; n0 = blockIdx.y * NQ;
        if (n0 > m0 + NB - 1) return;
        k0 = m0; k1 = J.Mp;
        A = J.X; B = J.X; C = J.T;
    } else {
        // group g of this level: blocks [a, a + h) left, [a + h, min(a + 2h, nb)) right; the tile lies in the block
        // (right rows, left columns)
        const int a = blockIdx.z * 2 * J.h, mid = a + J.h, end = min(a + 2 * J.h, nb);
        if (mid >= nb) return;
        m0 = (mid + blockIdx.x) * NB; n0 = a * NB + blockIdx.y * NQ;
        if (m0 >= end * NB || n0 >= mid * NB) return;
        if (J.mode == GEMM_TRTRI_T) {          // T[m, n] = sum_{k in left, k >= n0} L[m, k] X[k, n]
            A = J.L; B = J.X; C = J.T;
            k0 = n0; k1 = mid * NB;
        } else {                               // X[m, n] = -sum_{k in right, k <= m0 + 63} X[m, k] T[k, n]
            A = J.X; B = J.T; C = J.X;
            k0 = mid * NB; k1 = m0 + NB;
            negate = true;
        }
    }
    const int nk = (k1 - k0) / GK;
    const int lrow = tid >> 3, lch = tid & 7, vrow = tid >> 6, vch = tid & 63, trow = tid >> 5, tch = tid & 31;
    auto stage_load = [&](int s, int kt) {
        double* As = sm + (size_t)s * (GEMM_STAGE / 8);
        const int k = k0 + kt * GK;
        double* Bs;
        if (AT) {        // A^T: 16 rows (k) x 64 doubles (m): 512 chunks
            Bs = As + GK * G_LDAT;
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) cp_async16(As + (trow + 8 * hh) * G_LDAT + tch * 2, A + (size_t)(k + trow + 8 * hh) * ld + m0 + tch * 2);
        } else {         // 64 rows (m) x 16 doubles (k): 512 chunks
            Bs = As + NB * G_LDA;
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) cp_async16(As + (lrow + 32 * hh) * G_LDA + lch * 2, A + (size_t)(m0 + lrow + 32 * hh) * ld + k + lch * 2);
        }
        if (BT) {        // B^T: 128 rows (n) x 16 doubles (k): 1024 chunks
#pragma unroll
            for (int hh = 0; hh < 4; ++hh) cp_async16(Bs + (lrow + 32 * hh) * G_LDBT + lch * 2, B + (size_t)(n0 + lrow + 32 * hh) * ld + k + lch * 2);
        } else {         // 16 rows (k) x 128 doubles (n): 1024 chunks
#pragma unroll
            for (int hh = 0; hh < 4; ++hh) cp_async16(Bs + (vrow + 4 * hh) * G_LDB + vch * 2, B + (size_t)(k + vrow + 4 * hh) * ld + n0 + vch * 2);
        }
    };
    double acc[4][4][2] = {};
    for (int s = 0; s < GST - 1; ++s) {
        if (s < nk) stage_load(s, s);
        cp_async_commit();
    }
    for (int kt = 0; kt < nk; ++kt) {
        cp_async_wait<GST - 2>();
        __syncthreads();
        const int nx = kt + GST - 1;
        if (nx < nk) stage_load(nx % GST, nx);
        cp_async_commit();
        const double* As = sm + (size_t)(kt % GST) * (GEMM_STAGE / 8);
        const double* Bs = As + (AT ? GK * G_LDAT : NB * G_LDA);
        warp_mma<BT, GK / 4, AT>(acc, AT ? As + wm * 32 : As + wm * 32 * G_LDA, AT ? G_LDAT : G_LDA,
                                 BT ? Bs + wn * 32 * G_LDBT : Bs + wn * 32, BT ? G_LDBT : G_LDB, lane);
    }
    cp_async_wait<0>();
    const double sgn = negate ? -1.0 : 1.0;
#pragma unroll
    for (int mi = 0; mi < 4; ++mi)
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) {
            const int m = m0 + wm * 32 + mi * 8 + r, n = n0 + wn * 32 + ni * 8 + c * 2;
            double2* dst = reinterpret_cast<double2*>(C + (size_t)m * ld + n);
            if (subtract) { const double2 o = *dst; *dst = make_double2(o.x - acc[mi][ni][0], o.y - acc[mi][ni][1]); }
            else *dst = make_double2(sgn * acc[mi][ni][0], sgn * acc[mi][ni][1]);
        }
}

// first level of the recursive inverse (h = 1): pairs of 64 x 64 blocks, X_aa = Dinv_a, X_bb = Dinv_b,
// X_ba = -Dinv_b (L_ba Dinv_a); one CTA (128 threads) per pair of block rows (2g, 2g + 1)
__global__ void __launch_bounds__(128) gp_trtri_leaf_kernel(const double* L, const double* Dinv, int Mp, double* X) {
    extern __shared__ double sm[];
    double* As = sm;
    double* Bs = sm + NB * LDT;
    double* Ss = sm + 2 * NB * LDT;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5, wm = w >> 1, wn = w & 1;
    const int r = lane >> 2, c = lane & 3;
    const int a = 2 * blockIdx.x, b = a + 1;
    const double* Da = Dinv + (size_t)a * NB * NB;
    const double* Db = Dinv + (size_t)b * NB * NB;
    for (int e = tid; e < NB * NB; e += 128) {
        X[(size_t)(a * NB + e / NB) * Mp + a * NB + e % NB] = Da[e];
        X[(size_t)(b * NB + e / NB) * Mp + b * NB + e % NB] = Db[e];
        X[(size_t)(a * NB + e / NB) * Mp + b * NB + e % NB] = 0.0;        // above the diagonal
    }
    load_tile(As, L + (size_t)b * NB * Mp + (size_t)a * NB, Mp, tid, 128);
    load_tile(Bs, Da, NB, tid, 128);
    __syncthreads();
    double acc[4][4][2] = {};
    warp_mma<false, NB / 4>(acc, As + wm * 32 * LDT, LDT, Bs + wn * 32, LDT, lane);
#pragma unroll
    for (int mi = 0; mi < 4; ++mi)
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) {
            double* q = Ss + (wm * 32 + mi * 8 + r) * LDT + wn * 32 + ni * 8 + c * 2;
            q[0] = acc[mi][ni][0];
            q[1] = acc[mi][ni][1];
        }
    __syncthreads();
    load_tile(As, Db, NB, tid, 128);
    __syncthreads();
    double out[4][4][2] = {};
    warp_mma<false, NB / 4>(out, As + wm * 32 * LDT, LDT, Ss + wn * 32, LDT, lane);
    double* Xba = X + (size_t)b * NB * Mp + (size_t)a * NB;
#pragma unroll
    for (int mi = 0; mi < 4; ++mi)
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) {
            double* q = Xba + (size_t)(wm * 32 + mi * 8 + r) * Mp + wn * 32 + ni * 8 + c * 2;
            q[0] = -out[mi][ni][0];
            q[1] = -out[mi][ni][1];
        }
}

// partial gradient sums of one 64 x 64 tile of the lower triangle: part[tile][0] (log_amp),
// part[tile][1 + d] (log_scale_d).  Off-diagonal elements count twice (symmetry).
__global__ void __launch_bounds__(256) gp_grad_kernel(const double* Kinv, const double* alpha, const double* Xs, int M, int Mp,
                                                      int D, int nout, GpKern kp, double* part) {
    const int p = blockIdx.x;
    int bi = (int)((sqrt(8.0 * p + 1.0) - 1.0) * 0.5);
    while ((bi + 1) * (bi + 2) / 2 <= p) ++bi;
    while (bi * (bi + 1) / 2 > p) --bi;
    const int bj = p - bi * (bi + 1) / 2;
    double g[1 + GP_MAXD];
#pragma unroll
    for (int k = 0; k <= GP_MAXD; ++k) g[k] = 0.0;
    for (int e = threadIdx.x; e < NB * NB; e += 256) {
        const int i = bi * NB + e / NB, j = bj * NB + e % NB;
        if (i >= M || j > i) continue;
        double wgt = nout * Kinv[(size_t)i * Mp + j];
        for (int o = 0; o < nout; ++o) wgt -= alpha[(size_t)o * Mp + i] * alpha[(size_t)o * Mp + j];
        if (i != j) wgt *= 2.0;
        double d2 = 0.0, t2[GP_MAXD];
#pragma unroll
        for (int k = 0; k < GP_MAXD; ++k) {
            t2[k] = 0.0;
            if (k < D) { const double t = Xs[(size_t)i * D + k] - Xs[(size_t)j * D + k]; t2[k] = t * t; d2 += t2[k]; }
        }
        const double a = SQRT3 * sqrt(d2), ex = exp(-a);
        g[0] += wgt * kp.amp * (1.0 + a) * ex;
        const double f = wgt * 3.0 * kp.amp * ex;
#pragma unroll
        for (int k = 0; k < GP_MAXD; ++k) g[1 + k] += f * t2[k];
    }
    __shared__ double red[8][1 + GP_MAXD];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k <= GP_MAXD; ++k) {
        double v = g[k];
#pragma unroll
        for (int oo = 16; oo > 0; oo >>= 1) v += __shfl_xor_sync(0xffffffffu, v, oo);
        if (lane == 0) red[w][k] = v;
    }
    __syncthreads();
    if (threadIdx.x <= GP_MAXD) {
        double v = 0.0;
        for (int ww = 0; ww < 8; ++ww) v += red[ww][threadIdx.x];
        part[(size_t)p * (1 + GP_MAXD) + threadIdx.x] = v;
    }
}

// out[0] = loss, out[1 + k] = d loss / d theta_k; one block, fixed summation order
__global__ void __launch_bounds__(256) gp_nlml_finish_kernel(const double* L, const double* alpha, const double* y, int M,
                                                             int Mp, int nout, const double* part, int ntile, double* out) {
    __shared__ double red[256];
    const int tid = threadIdx.x;
    double quad = 0.0, logdet = 0.0;
    for (int i = tid; i < M; i += 256) {
        logdet += log(L[(size_t)i * Mp + i]);
        for (int o = 0; o < nout; ++o) quad += y[(size_t)o * Mp + i] * alpha[(size_t)o * Mp + i];
    }
    double vals[2 + 1 + GP_MAXD];
    vals[0] = quad; vals[1] = logdet;
    for (int k = 0; k <= GP_MAXD; ++k) {
        double v = 0.0;
        for (int t = tid; t < ntile; t += 256) v += part[(size_t)t * (1 + GP_MAXD) + k];
        vals[2 + k] = v;
    }
    for (int q = 0; q < 3 + GP_MAXD; ++q) {
        red[tid] = vals[q];
        __syncthreads();
        for (int s = 128; s > 0; s >>= 1) {
            if (tid < s) red[tid] += red[tid + s];
            __syncthreads();
        }
        if (tid == 0) vals[q] = red[0];
        __syncthreads();
    }
    if (tid == 0) {
        out[0] = 0.5 * vals[0] + nout * (vals[1] + 0.5 * M * 1.8378770664093453);      // log(2 pi)
        for (int k = 0; k <= GP_MAXD; ++k) out[1 + k] = 0.5 * vals[2 + k];
    }
}

// =====================================================================================
// Host side
// =====================================================================================
static size_t al(size_t x) { return (x + 255) / 256 * 256; }

int GpModel::var_slots(int sms) { return VAR_CTAS_PER_SM * sms; }

size_t GpModel::workspace_bytes(int Mmax, int D, int nout, int nslots) {
    const size_t Mp = pad_m(Mmax);
    size_t b = 256;
    b += al(Mp * D * 8);                 // Xs
    b += al(Mp * Mp * 8);                // L
    b += al(Mp / NB * NB * NB * 8);      // Dinv
    b += al(Mp * 2 * 8);                 // alpha (column-major, 2 columns)
    b += al(Mp * 2 * 8);                 // copy of the right-hand sides (training loss)
    b += al((size_t)nslots * Mp * NQ * 8);   // v panels of the variance kernel
    b += al(65536 * 8) + al(4096 * 8) + al(4096 * 8) + 256;   // partial maxima, argmax parts, scalars
    b += al(Mp * 5 * 8) + al((2 * Mp + 1) * 4);               // tables of the one-input mean (gp_mean1d.h)
    b += al(Mp * 7 * 8);                                      // tables of the one-input variance (gp_var1d.h)
    (void)nout;
    return b;
}

int GpModel::init(int Mmax_, int D_, int nout_, void* ws, size_t bytes) {
    if (Mmax_ < 1 || D_ < 1 || D_ > GP_MAXD || nout_ < 1 || nout_ > 2) { set_error("bad GP dimensions (D <= 8, n_out <= 2)"); return GPF_E_INVALID; }
    Mmax = Mmax_; D = D_; nout = nout_; M = 0; Mp = 0;
    int dev = 0, sms = 148;
    GPF_CUDA(cudaGetDevice(&dev));
    GPF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    nslots = VAR_CTAS_PER_SM * sms;     // one v panel per resident CTA of the variance kernel
    {
        // the opt-in to > 48 KB of dynamic shared memory is a per-device function attribute: set it for the device
        // of every model (cheap), not once per process
        const int sm2 = (int)((size_t)2 * NB * LDT * 8), sm3 = (int)((size_t)3 * NB * LDT * 8);
        GPF_CUDA(cudaFuncSetAttribute(gp_potrf_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)POTRF_SMEM));
        GPF_CUDA(cudaFuncSetAttribute(gp_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, sm2));
        GPF_CUDA(cudaFuncSetAttribute(gp_syrk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, sm2));
        GPF_CUDA(cudaFuncSetAttribute(gp_var_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)VAR_SMEM));
        GPF_CUDA(cudaFuncSetAttribute(gp_trtri_leaf_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, sm3));
        GPF_CUDA(cudaFuncSetAttribute(gp_gemm_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GEMM_SMEM));
        GPF_CUDA(cudaFuncSetAttribute(gp_gemm_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GEMM_SMEM));
        GPF_CUDA(cudaFuncSetAttribute(gp_gemm_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GEMM_SMEM));
    }
    if (bytes < workspace_bytes(Mmax, D, nout, nslots)) { set_error("GP workspace too small"); return GPF_E_INVALID; }
    const size_t Mpm = pad_m(Mmax);
    unsigned char* p = reinterpret_cast<unsigned char*>(al(reinterpret_cast<size_t>(ws)));
    Xs = reinterpret_cast<double*>(p); p += al(Mpm * D * 8);
    L = reinterpret_cast<double*>(p); p += al(Mpm * Mpm * 8);
    Dinv = reinterpret_cast<double*>(p); p += al(Mpm / NB * NB * NB * 8);
    alpha = reinterpret_cast<double*>(p); p += al(Mpm * 2 * 8);
    ycopy = reinterpret_cast<double*>(p); p += al(Mpm * 2 * 8);
    scratch = reinterpret_cast<double*>(p); p += al((size_t)nslots * Mpm * NQ * 8);
    gpart = reinterpret_cast<double*>(p); p += al(65536 * 8);
    vpart = reinterpret_cast<double*>(p); p += al(4096 * 8);
    ipart = reinterpret_cast<long long*>(p); p += al(4096 * 8);
    scal = reinterpret_cast<double*>(p); p += 256;
    t1d = reinterpret_cast<double*>(p); p += al(Mpm * 5 * 8);
    t1d_i = reinterpret_cast<int*>(p); p += al((2 * Mpm + 1) * 4);
    v1d = reinterpret_cast<double*>(p);
    t1d_valid = false;
    v1d_valid = false;
    return GPF_OK;
}

int GpModel::factor(const double* Xtrain, int M_, double log_amp, const double* log_scale, double yerr, const double* Y,
                    cudaStream_t s) {
    if (M_ < 1 || M_ > Mmax) { set_error("training-set size outside [1, M_max]"); return GPF_E_INVALID; }
    M = M_;
    Mp = (int)pad_m(M);
    kp.D = D;
    kp.amp = exp(log_amp);
    for (int k = 0; k < GP_MAXD; ++k) kp.s[k] = k < D ? exp(-log_scale[k]) : 0.0;
    // pre-scaled training inputs and right-hand sides (host -> device)
    std::vector<double> xs((size_t)Mp * D, 0.0), y((size_t)Mp * 2, 0.0);
    for (int i = 0; i < M; ++i)
        for (int k = 0; k < D; ++k) xs[(size_t)i * D + k] = Xtrain[(size_t)i * D + k] * kp.s[k];
    for (int i = 0; i < M; ++i)
        for (int o = 0; o < nout; ++o) y[(size_t)o * Mp + i] = Y[(size_t)i * nout + o];
    t1d_valid = false;
    v1d_valid = false;
    yerr_last = yerr;
    if (D == 1) u_host.assign(xs.begin(), xs.begin() + M);
    GPF_CUDA(cudaMemcpyAsync(Xs, xs.data(), xs.size() * 8, cudaMemcpyHostToDevice, s));
    GPF_CUDA(cudaMemcpyAsync(alpha, y.data(), (size_t)Mp * 2 * 8, cudaMemcpyHostToDevice, s));
    GPF_CUDA(cudaMemcpyAsync(ycopy, y.data(), (size_t)Mp * 2 * 8, cudaMemcpyHostToDevice, s));
    GPF_CUDA(cudaStreamSynchronize(s));       // the host vectors go out of scope
    int* info = reinterpret_cast<int*>(scal + 8);
    GPF_CUDA(cudaMemsetAsync(info, 0, sizeof(int), s));
    gp_build_K_kernel<<<dim3((Mp + 255) / 256, Mp), 256, 0, s>>>(Xs, M, Mp, D, kp.amp, yerr * yerr, L);
    g_launches++;
    const int nb = Mp / NB;
    const size_t sm2 = (size_t)2 * NB * LDT * 8;
    const size_t smp = POTRF_SMEM;
    // Blocked right-looking Cholesky with a 256-wide outer panel (4 blocks of 64).  Inside the panel: diagonal block,
    // panel solve over all rows below, update of the panel's remaining columns.  Everything right of the panel is
    // updated once per panel by the GEMM engine with k-depth 256 -- first the strip that the NEXT panel consists of
    // (on the caller's stream, which then factors that panel), the rest concurrently on a second stream: the serial
    // chain of small kernels (diagonal blocks: 6 ms per factorisation) runs under the large update, not after it.
    constexpr int KO = 4;
    if (!side) {
        GPF_CUDA(cudaStreamCreateWithFlags(&side, cudaStreamNonBlocking));
        GPF_CUDA(cudaEventCreateWithFlags(&ev_strip, cudaEventDisableTiming));
        GPF_CUDA(cudaEventCreateWithFlags(&ev_rest, cudaEventDisableTiming));
    }
    bool rest_pending = false;
    for (int K0 = 0; K0 < nb; K0 += KO) {
        const int kend = std::min(K0 + KO, nb);
        for (int k = K0; k < kend; ++k) {
            gp_potrf_diag_kernel<<<1, 256, smp, s>>>(L, Mp, k, Dinv, info);
            g_launches++;
            const int rem = nb - 1 - k;
            if (rem > 0) {
                gp_panel_kernel<<<rem, 128, sm2, s>>>(L, Mp, k, Dinv);
                g_launches++;
                const int ncol = kend - 1 - k;
                if (ncol > 0) { gp_syrk_kernel<<<dim3(rem, ncol), 128, sm2, s>>>(L, Mp, k); g_launches++; }
            }
        }
        if (kend >= nb) break;
        // trailing update with this panel: rows and columns from block kend on
        if (rest_pending) { GPF_CUDA(cudaStreamWaitEvent(s, ev_rest, 0)); rest_pending = false; }   // columns it is about to change
        GemmJob J;
        J.mode = GEMM_SYRK; J.Mp = Mp; J.h = 0; J.L = L; J.X = L; J.T = L; J.kb0 = K0; J.kb1 = kend;
        const int rows = nb - kend, ctiles = (rows + 1) / 2, strip = std::min(KO / 2, ctiles);
        if (ctiles > strip) GPF_CUDA(cudaEventRecord(ev_strip, s));    // the panel is final: the side stream may read it
        J.ct0 = 0;
        gp_gemm_kernel<false, true><<<dim3(rows, strip), 256, GEMM_SMEM, s>>>(J);
        g_launches++;
        if (ctiles > strip) {
            GPF_CUDA(cudaStreamWaitEvent(side, ev_strip, 0));
            J.ct0 = strip;
            gp_gemm_kernel<false, true><<<dim3(rows, ctiles - strip), 256, GEMM_SMEM, side>>>(J);
            g_launches++;
            GPF_CUDA(cudaEventRecord(ev_rest, side));
            rest_pending = true;
        }
    }
    if (rest_pending) GPF_CUDA(cudaStreamWaitEvent(s, ev_rest, 0));
    // alpha: the right-hand sides are in `alpha`; the variance kernel's scratch serves as the z vector
    double* zvec = scratch;
    for (int k = 0; k < nb; ++k) { gp_trsv_fwd_kernel<<<nb - k, 256, 0, s>>>(L, Mp, k, Dinv, alpha, zvec, nout); g_launches++; }
    for (int k = nb - 1; k >= 0; --k) { gp_trsv_bwd_kernel<<<k + 1, 256, 0, s>>>(L, Mp, k, Dinv, zvec, alpha, nout); g_launches++; }
    GPF_CUDA(cudaGetLastError());
    int h = 0;
    GPF_CUDA(cudaMemcpyAsync(&h, info, sizeof(int), cudaMemcpyDeviceToHost, s));
    GPF_CUDA(cudaStreamSynchronize(s));
    if (h) { set_error("K + yerr^2 I is not positive definite at row " + std::to_string(h - 1)); M = 0; return GPF_E_INVALID; }
    return GPF_OK;
}

size_t GpModel::train_workspace_bytes(int Mmax) {
    const size_t Mp = pad_m(Mmax), nb = Mp / NB;
    return 256 + 2 * al(Mp * Mp * 8) + al(nb * (nb + 1) / 2 * (1 + GP_MAXD) * 8) + al((2 + GP_MAXD) * 8);
}

int GpModel::nlml(void* work, size_t bytes, double* loss, double* grad, cudaStream_t s) {
    if (M < 1) { set_error("no factorisation"); return GPF_E_STATE; }
    if (!work || bytes < train_workspace_bytes(M)) { set_error("training workspace too small"); return GPF_E_INVALID; }
    const size_t Mpp = Mp;
    const int nb = Mp / NB, ntile = nb * (nb + 1) / 2;
    unsigned char* p = reinterpret_cast<unsigned char*>(al(reinterpret_cast<size_t>(work)));
    double* X = reinterpret_cast<double*>(p); p += al(Mpp * Mpp * 8);
    double* Kinv = reinterpret_cast<double*>(p); p += al(Mpp * Mpp * 8);
    double* part = reinterpret_cast<double*>(p); p += al((size_t)ntile * (1 + GP_MAXD) * 8);
    double* out = reinterpret_cast<double*>(p);
    const size_t sm3 = (size_t)3 * NB * LDT * 8, sm2 = (size_t)2 * NB * LDT * 8;
    // X = L^-1 by recursive doubling over the block rows (nb is even: Mp is a multiple of 128):
    //   level 0 pairs of 64-blocks (leaf kernel); level h = 2, 4, ...: X21 = -X22 (L21 X11), two batched GEMMs
    //   whose tiles all run at once (the block-column kernel this replaces was one serial chain per column:
    //   141 of the 160 ms of an objective evaluation at M = 10^4, DMMA pipe 18 % busy: profiles/r2a)
    gp_trtri_leaf_kernel<<<nb / 2, 128, sm3, s>>>(L, Dinv, Mp, X);
    g_launches += 1;
    for (int h = 2; h < nb; h *= 2) {
        GemmJob J;
        J.Mp = Mp; J.h = h; J.L = L; J.X = X; J.T = Kinv;          // Kinv's storage is free until the last step
        const dim3 grid(h, h / 2, (nb + 2 * h - 1) / (2 * h));      // (right block rows, left 128-column tiles, groups)
        J.mode = GEMM_TRTRI_T;
        gp_gemm_kernel<false, false><<<grid, 256, GEMM_SMEM, s>>>(J);
        J.mode = GEMM_TRTRI_X;
        gp_gemm_kernel<false, false><<<grid, 256, GEMM_SMEM, s>>>(J);
        g_launches += 2;
    }
    {
        GemmJob J;
        J.mode = GEMM_LAUUM; J.Mp = Mp; J.h = 0; J.L = L; J.X = X; J.T = Kinv;
        gp_gemm_kernel<true, false><<<dim3(nb, Mp / NQ), 256, GEMM_SMEM, s>>>(J);
        g_launches += 1;
    }
    gp_grad_kernel<<<ntile, 256, 0, s>>>(Kinv, alpha, Xs, M, Mp, D, nout, kp, part);
    gp_nlml_finish_kernel<<<1, 256, 0, s>>>(L, alpha, ycopy, M, Mp, nout, part, ntile, out);
    g_launches += 2;
    GPF_CUDA(cudaGetLastError());
    double h[2 + GP_MAXD];
    GPF_CUDA(cudaMemcpyAsync(h, out, sizeof(h), cudaMemcpyDeviceToHost, s));
    GPF_CUDA(cudaStreamSynchronize(s));
    if (loss) *loss = h[0];
    if (grad) for (int k = 0; k <= D; ++k) grad[k] = h[1 + k];
    return GPF_OK;
}

int GpModel::get_factor(double* L_out, double* alpha_out, cudaStream_t s) {
    if (M < 1) { set_error("no factorisation"); return GPF_E_STATE; }
    if (L_out) {
        // the strict upper triangle of the device array still holds K (never read): report zeros
        GPF_CUDA(cudaMemcpy2DAsync(L_out, (size_t)M * 8, L, (size_t)Mp * 8, (size_t)M * 8, M, cudaMemcpyDeviceToHost, s));
        GPF_CUDA(cudaStreamSynchronize(s));
        for (int i = 0; i < M; ++i)
            for (int j = i + 1; j < M; ++j) L_out[(size_t)i * M + j] = 0.0;
    }
    if (alpha_out) {
        std::vector<double> a((size_t)Mp * 2);
        GPF_CUDA(cudaMemcpyAsync(a.data(), alpha, a.size() * 8, cudaMemcpyDeviceToHost, s));
        GPF_CUDA(cudaStreamSynchronize(s));
        for (int i = 0; i < M; ++i)
            for (int o = 0; o < nout; ++o) alpha_out[(size_t)i * nout + o] = a[(size_t)o * Mp + i];
    }
    GPF_CUDA(cudaStreamSynchronize(s));
    return GPF_OK;
}

// ---- one-input models: O(log M) mean per query (gp_mean1d.h) ---------------------------------------------
template <bool GRAD>
__global__ void __launch_bounds__(256) gp_mean1d_kernel(Mean1D T, GpKern kp, const double* Xtest, long long N, long long off,
                                                        long long cnt, double* mean, double* gmax_part) {
    __shared__ double sred[8];
    const long long i = off + (long long)blockIdx.x * 256 + threadIdx.x;
    double g = -INFINITY;
    if (i < off + cnt) {
        double m, gs;
        mean1d_eval(T, Xtest[i] * kp.s[0], &m, &gs);
        mean[i] = kp.amp * m;
        g = -3.0 * kp.amp * kp.s[0] * gs;
    }
    if (!GRAD) return;
#pragma unroll
    for (int oo = 16; oo > 0; oo >>= 1) g = fmax(g, __shfl_xor_sync(0xffffffffu, g, oo));
    if ((threadIdx.x & 31) == 0) sred[threadIdx.x >> 5] = g;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int ww = 1; ww < 8; ++ww) g = fmax(g, sred[ww]);
        gmax_part[blockIdx.x] = g;
    }
}

// tables of the last factorisation: alpha comes back from the device (M doubles), the running sums are built on
// the host (M log M, long double) and go up again (5 M doubles)
int GpModel::build_mean1d(cudaStream_t s) {
    std::vector<double> a(M);
    GPF_CUDA(cudaMemcpyAsync(a.data(), alpha, (size_t)M * 8, cudaMemcpyDeviceToHost, s));
    GPF_CUDA(cudaStreamSynchronize(s));
    Mean1DHost H;
    H.build(u_host.data(), a.data(), M);
    const size_t Mpm = pad_m(Mmax);
    double* d[5] = {t1d, t1d + Mpm, t1d + 2 * Mpm, t1d + 3 * Mpm, t1d + 4 * Mpm};
    const std::vector<double>* h[5] = {&H.u, &H.A, &H.B, &H.C, &H.D};
    for (int k = 0; k < 5; ++k) GPF_CUDA(cudaMemcpyAsync(d[k], h[k]->data(), (size_t)M * 8, cudaMemcpyHostToDevice, s));
    GPF_CUDA(cudaMemcpyAsync(t1d_i, H.blk.data(), (size_t)M * 4, cudaMemcpyHostToDevice, s));
    GPF_CUDA(cudaMemcpyAsync(t1d_i + Mpm, H.bstart.data(), H.bstart.size() * 4, cudaMemcpyHostToDevice, s));
    GPF_CUDA(cudaStreamSynchronize(s));                  // the host vectors go out of scope
    t1d_view.M = M; t1d_view.nb = (int)H.bstart.size() - 1;
    t1d_view.u = d[0]; t1d_view.A = d[1]; t1d_view.B = d[2]; t1d_view.C = d[3]; t1d_view.D = d[4];
    t1d_view.blk = t1d_i; t1d_view.bstart = t1d_i + Mpm;
    t1d_valid = true;
    return GPF_OK;
}

int GpModel::mean(const double* Xtest, long long N, double* mean_out, double* dmax_host, cudaStream_t s) {
    if (M < 1) { set_error("no factorisation"); return GPF_E_STATE; }
    if (N < 1) return GPF_OK;
    static const bool use1d = [] { const char* e = getenv("GPF_GP_MEAN1D"); return !e || atoi(e) != 0; }();
    if (D == 1 && nout == 1 && use1d) {
        if (!t1d_valid) { const int rc = build_mean1d(s); if (rc) return rc; }
        double gmax = -INFINITY;
        const long long chunk = 65536LL * 256;
        for (long long off = 0; off < N; off += chunk) {
            const long long cnt = std::min<long long>(chunk, N - off);
            const int nbk = (int)((cnt + 255) / 256);
            if (dmax_host) gp_mean1d_kernel<true><<<nbk, 256, 0, s>>>(t1d_view, kp, Xtest, N, off, cnt, mean_out, gpart);
            else gp_mean1d_kernel<false><<<nbk, 256, 0, s>>>(t1d_view, kp, Xtest, N, off, cnt, mean_out, gpart);
            g_launches += 1;
            if (dmax_host) {
                gp_max_reduce_kernel<<<1, 256, 0, s>>>(gpart, nbk, scal);
                g_launches += 1;
                double g;
                GPF_CUDA(cudaMemcpyAsync(&g, scal, 8, cudaMemcpyDeviceToHost, s));
                GPF_CUDA(cudaStreamSynchronize(s));
                gmax = std::fmax(gmax, g);
            }
        }
        GPF_CUDA(cudaGetLastError());
        if (dmax_host) *dmax_host = gmax;
        return GPF_OK;
    }
    // at most 65536 partial maxima per launch
    double gmax = -INFINITY;
    const long long chunk = 65536LL * 512;                  // 512 queries per block (two per thread)
    for (long long off = 0; off < N; off += chunk) {
        const long long cnt = std::min<long long>(chunk, N - off);
        const int nbk = (int)((cnt + 511) / 512);
        const bool grad = dmax_host != nullptr;
        switch (D) {
#define GP_MEAN_LAUNCH(DD, NO, GR) do { if (fast_arith) gp_mean_kernel<DD, NO, GR, true><<<nbk, 256, 0, s>>>(Xs, alpha, M, Mp, nout, kp, Xtest, N, off, cnt, mean_out, gpart); \
                                        else gp_mean_kernel<DD, NO, GR, false><<<nbk, 256, 0, s>>>(Xs, alpha, M, Mp, nout, kp, Xtest, N, off, cnt, mean_out, gpart); } while (0)
#define GP_MEAN_CASE(DD) case DD:                                                     \
            if (grad) { if (nout > 1) GP_MEAN_LAUNCH(DD, 2, true); else GP_MEAN_LAUNCH(DD, 1, true); }   \
            else { if (nout > 1) GP_MEAN_LAUNCH(DD, 2, false); else GP_MEAN_LAUNCH(DD, 1, false); }      \
            break;
            GP_MEAN_CASE(1) GP_MEAN_CASE(2) GP_MEAN_CASE(3) GP_MEAN_CASE(4) GP_MEAN_CASE(5) GP_MEAN_CASE(6) GP_MEAN_CASE(7) GP_MEAN_CASE(8)
#undef GP_MEAN_CASE
#undef GP_MEAN_LAUNCH
        }
        g_launches += 1;
        if (dmax_host) {
            gp_max_reduce_kernel<<<1, 256, 0, s>>>(gpart, nbk, scal);
            g_launches += 1;
            double g;
            GPF_CUDA(cudaMemcpyAsync(&g, scal, 8, cudaMemcpyDeviceToHost, s));
            GPF_CUDA(cudaStreamSynchronize(s));
            gmax = std::fmax(gmax, g);
        }
    }
    GPF_CUDA(cudaGetLastError());
    if (dmax_host) *dmax_host = gmax;
    return GPF_OK;
}

// ---- one-input models: O(log M) variance per query (gp_var1d.h) ------------------------------------------
// grid-stride over the queries; per block the largest variance and the index of its first occurrence (NaN wins,
// like np.argmax), folded by gp_argmax_reduce_kernel exactly as the dense kernel's per-CTA records are
__global__ void __launch_bounds__(256) gp_var1d_kernel(Var1D T, GpKern kp, const double* Xtest, long long N,
                                                       const double* noise, double noise_scalar, double* var,
                                                       double* vpart, long long* ipart) {
    __shared__ double sv[256];
    __shared__ long long si[256];
    double b = -INFINITY;
    long long bi = 0x7fffffffffffffffLL;
    bool nanv = false;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < N; i += (long long)gridDim.x * 256) {
        const double v = var1d_eval(T, Xtest[i] * kp.s[0]) + (noise ? noise[i] : noise_scalar);
        var[i] = v;
        if (v != v) { if (!nanv) { nanv = true; b = v; bi = i; } continue; }     // i ascends: the first NaN stays
        if (!nanv && v > b) { b = v; bi = i; }                                   // strict: the first maximum stays
    }
    sv[threadIdx.x] = b;
    si[threadIdx.x] = bi;
    __syncthreads();
    if (threadIdx.x != 0) return;
    nanv = false; b = -INFINITY; bi = 0x7fffffffffffffffLL;
    for (int t = 0; t < 256; ++t) {
        const double v = sv[t];
        if (v != v) { if (!nanv || si[t] < bi) { nanv = true; b = v; bi = si[t]; } continue; }
        if (nanv) continue;
        if (v > b || (v == b && si[t] < bi)) { b = v; bi = si[t]; }
    }
    vpart[blockIdx.x] = b;
    ipart[blockIdx.x] = bi;
}

int GpModel::build_var1d(cudaStream_t s) {
    Var1DHost H;
    H.build(u_host.data(), M, kp.amp, yerr_last);
    const size_t Mpm = pad_m(Mmax);
    GPF_CUDA(cudaMemcpyAsync(v1d, H.u.data(), (size_t)M * 8, cudaMemcpyHostToDevice, s));
    GPF_CUDA(cudaMemcpyAsync(v1d + Mpm, H.Pf.data(), (size_t)3 * M * 8, cudaMemcpyHostToDevice, s));
    GPF_CUDA(cudaMemcpyAsync(v1d + 4 * Mpm, H.Pb.data(), (size_t)3 * M * 8, cudaMemcpyHostToDevice, s));
    GPF_CUDA(cudaStreamSynchronize(s));                  // the host vectors go out of scope
    v1d_view.M = M; v1d_view.amp = kp.amp;
    v1d_view.u = v1d; v1d_view.Pf = v1d + Mpm; v1d_view.Pb = v1d + 4 * Mpm;
    v1d_valid = true;
    return GPF_OK;
}

int GpModel::mean_var(const double* Xtest, long long N, const double* noise, double noise_scalar, double* mean_out,
                      double* var_out, double* vmax_host, long long* imax_host, cudaStream_t s) {
    int rc = mean(Xtest, N, mean_out, nullptr, s);
    if (rc) return rc;
    if (N < 1) return GPF_OK;
    // one-input models: the O(log M) two-filter form (GPF_GP_VAR1D=0 restores the dense solve)
    static const bool use1d = [] { const char* e = getenv("GPF_GP_VAR1D"); return !e || atoi(e) != 0; }();
    int grid;
    if (D == 1 && use1d) {
        if (!v1d_valid && (rc = build_var1d(s))) return rc;
        grid = (int)std::min<long long>((N + 255) / 256, 4096);
        gp_var1d_kernel<<<grid, 256, 0, s>>>(v1d_view, kp, Xtest, N, noise, noise_scalar, var_out, vpart, ipart);
    } else {
        const long long ntiles = (N + NQ - 1) / NQ;
        grid = (int)std::min<long long>(ntiles, nslots);
        gp_var_kernel<<<grid, 256, VAR_SMEM, s>>>(L, Dinv, Xs, M, Mp, kp, Xtest, N, noise, noise_scalar, var_out, scratch, vpart,
                                                  ipart);
    }
    gp_argmax_reduce_kernel<<<1, 32, 0, s>>>(vpart, ipart, grid, scal + 1, reinterpret_cast<long long*>(scal + 2));
    g_launches += 2;
    GPF_CUDA(cudaGetLastError());
    if (vmax_host || imax_host) {
        double h[2];
        GPF_CUDA(cudaMemcpyAsync(h, scal + 1, 16, cudaMemcpyDeviceToHost, s));
        GPF_CUDA(cudaStreamSynchronize(s));
        if (vmax_host) *vmax_host = h[0];
        if (imax_host) memcpy(imax_host, &h[1], 8);
    }
    return GPF_OK;
}

int launch_gp_features(FvSolver& S, const double* x_scale7, const int* dims, int D, double* X, cudaStream_t s, int rel) {
    if (D < 1 || D > GP_MAXD) { set_error("1 <= D <= 8"); return GPF_E_INVALID; }
    if (rel < 0 || rel > 2) { set_error("rel must be 0, 1 or 2"); return GPF_E_INVALID; }
    FeatArgs fa;
    fa.D = D;
    for (int k = 0; k < D; ++k) {
        if (dims[k] < 0 || dims[k] > 6) { set_error("active dimension outside [0, 6]"); return GPF_E_INVALID; }
        fa.dims[k] = dims[k];
        fa.scale[k] = x_scale7[dims[k]];
    }
    int cur;
    int rc = S.read_cur(s, &cur);
    if (rc) return rc;
    const int Nx2 = S.cfg.Nx + 2, Ny2 = S.cfg.Ny + 2;
    const size_t np = (size_t)Nx2 * Ny2;
    gp_features_kernel<<<(int)((np + 255) / 256), 256, 0, s>>>(S.qbuf[(cur + rel) % 3], S.topo, S.extra, S.P.L.plane, S.P.L.pitch, Nx2, Ny2,
                                                             fa, X);
    g_launches++;
    GPF_CUDA(cudaGetLastError());
    return GPF_OK;
}

}  // namespace gpf
