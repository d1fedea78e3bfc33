// Predictive mean of a ONE-input Matern-3/2 GP in O(log M) per query instead of O(M).
//
// gp.py:525-535 evaluates  mean(x) = amp * sum_j alpha_j (1 + c r_j) exp(-c r_j),  r_j = |u - u_j|,  c = sqrt(3),
// u = x * exp(-log_scale).  In one dimension the kernel is semiseparable: for the training points left of the query
// (u_j <= u), relative to a reference point rho,
//     sum_j alpha_j (1 + c (u - u_j)) e^{-c (u - u_j)} = e^{-c (u - rho)} [ (1 + c (u - rho)) A - c B ],
//     A = sum_j alpha_j e^{c (u_j - rho)},   B = sum_j alpha_j (u_j - rho) e^{c (u_j - rho)},
// and mirrored for the points right of it.  The sorted training points are cut into blocks no wider than W / c
// (so that no exponent inside a block exceeds W and the cancellation in (1 + c d) A - c B stays below a factor W);
// per point the tables hold the running sums of its block from the block's start (A, B; reference = the block's
// largest point) and to its end (C, D; reference = its smallest point).  A query bisects for its position, takes
// the partial sums of its own block and then whole blocks outwards until e^{-c d} underflows exactly as exp(-r)
// does in the direct sum.  The slope sum of stress.py:522-537, sum_j alpha_j (u - u_j) e^{-c r_j}, comes from the
// same tables.  Same arithmetic on host and device (this header): the host instantiation is what the CPU tests
// compare with the oracle; the device kernel calls the same function per query.
#pragma once
#include <math.h>
#include <algorithm>
#include <numeric>
#include <vector>

#ifdef __CUDACC__
#define GPF_HD __host__ __device__
#else
#define GPF_HD
#endif

namespace gpf {

struct Mean1D {
    int M, nb;
    const double* u;       // [M]  sorted scaled training inputs
    const double* A;       // [M]  prefix sums of the point's block (reference: block maximum)
    const double* B;
    const double* C;       // [M]  suffix sums of the point's block (reference: block minimum)
    const double* D;
    const int* blk;        // [M]  block of a point
    const int* bstart;     // [nb + 1]
};

constexpr double MEAN1D_W = 8.0;          // widest exponent inside a block
constexpr double MEAN1D_CUT = 750.0;      // exp(-750) == 0 in binary64

// m = sum_j alpha_j (1 + c r_j) e^{-c r_j},   g = sum_j alpha_j (u - u_j) e^{-c r_j}
GPF_HD inline void mean1d_eval(const Mean1D& T, double uq, double* m_out, double* g_out) {
    const double c = 1.7320508075688772;
    int lo = 0, hi = T.M;                               // k = #{j : u_j <= uq}
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (T.u[mid] <= uq) lo = mid + 1; else hi = mid;
    }
    const int k = lo;
    double m = 0.0, g = 0.0;
    if (k > 0) {
        int idx = k - 1, b = T.blk[idx];
        while (true) {
            const double d = uq - T.u[T.bstart[b + 1] - 1];
            if (c * d > MEAN1D_CUT) break;
            const double ex = exp(-c * d);
            m += ex * ((1.0 + c * d) * T.A[idx] - c * T.B[idx]);
            g += ex * (d * T.A[idx] - T.B[idx]);
            if (--b < 0) break;
            idx = T.bstart[b + 1] - 1;
        }
    }
    if (k < T.M) {
        int idx = k, b = T.blk[idx];
        while (true) {
            const double d = T.u[T.bstart[b]] - uq;
            if (c * d > MEAN1D_CUT) break;
            const double ex = exp(-c * d);
            m += ex * ((1.0 + c * d) * T.C[idx] + c * T.D[idx]);
            g -= ex * (d * T.C[idx] + T.D[idx]);
            if (++b >= T.nb) break;
            idx = T.bstart[b];
        }
    }
    *m_out = m;
    *g_out = g;
}

// host-side tables; running sums in long double so that a table entry carries one rounding
struct Mean1DHost {
    std::vector<double> u, A, B, C, D;
    std::vector<int> blk, bstart;
    void build(const double* u_in, const double* alpha, int M) {
        const double c = 1.7320508075688772;
        std::vector<int> order(M);
        std::iota(order.begin(), order.end(), 0);
        std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return u_in[a] < u_in[b]; });
        u.resize(M); A.resize(M); B.resize(M); C.resize(M); D.resize(M); blk.resize(M);
        std::vector<double> al(M);
        for (int j = 0; j < M; ++j) { u[j] = u_in[order[j]]; al[j] = alpha[order[j]]; }
        bstart.assign(1, 0);
        for (int j = 0; j < M; ++j) {
            if (j > 0 && c * (u[j] - u[bstart.back()]) > MEAN1D_W) bstart.push_back(j);
            blk[j] = (int)bstart.size() - 1;
        }
        bstart.push_back(M);
        const int nb = (int)bstart.size() - 1;
        for (int b = 0; b < nb; ++b) {
            const int lo = bstart[b], hi = bstart[b + 1];
            const double rho = u[hi - 1], sig = u[lo];
            long double a = 0, bb = 0;
            for (int j = lo; j < hi; ++j) {
                const long double e = expl((long double)c * ((long double)u[j] - rho));
                a += (long double)al[j] * e;
                bb += (long double)al[j] * ((long double)u[j] - rho) * e;
                A[j] = (double)a; B[j] = (double)bb;
            }
            long double cc = 0, dd = 0;
            for (int j = hi - 1; j >= lo; --j) {
                const long double e = expl(-(long double)c * ((long double)u[j] - sig));
                cc += (long double)al[j] * e;
                dd += (long double)al[j] * ((long double)u[j] - sig) * e;
                C[j] = (double)cc; D[j] = (double)dd;
            }
        }
    }
    Mean1D view() const {
        Mean1D T;
        T.M = (int)u.size(); T.nb = (int)bstart.size() - 1;
        T.u = u.data(); T.A = A.data(); T.B = B.data(); T.C = C.data(); T.D = D.data();
        T.blk = blk.data(); T.bstart = bstart.data();
        return T;
    }
};

}  // namespace gpf
