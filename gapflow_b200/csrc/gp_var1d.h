// Predictive variance of a ONE-input Matern-3/2 GP in O(log M) per query instead of O(M^2).
//
// gp.py:509-522 evaluates  var(x) = k(x, x) + noise - || L^-1 k(X, x) ||^2  (M^2 flops per query).  In one
// dimension the Matern-3/2 process is the stationary solution of a two-state linear SDE (state z = (f, f')):
//     transition over a distance d >= 0:   A(d) = e^{-c d} [[1 + c d, d], [-c^2 d, 1 - c d]],   c = sqrt(3)
//     stationary covariance:               Pinf = diag(amp, c^2 amp)
//     prediction of a covariance P:        Pinf + A(d) (P - Pinf) A(d)^T
// and the GP posterior given noisy observations (variance yerr^2) at the sorted training points is what a Kalman
// filter computes.  The process is reversible (time reversal flips the sign of f'), so the same filter run from
// the right gives the posterior given the points to the right.  Tables: per training point the filtered
// covariance given the points up to it (Pf) and given the points from it on (Pb, in the mirrored coordinates).
// A query between points k and k+1 predicts Pf_k forward and Pb_{k+1} backward to its position and combines the
// two posteriors, which share the prior, in information form:
//     P^-1 = Pf*^-1 + Pb*^-1 - Pinf^-1,        var = P[0][0]  (+ noise)
// 2 x 2 symmetric algebra and two exponentials per query.  No cancellation of amp against the quadratic form:
// the result is at least as accurate as the reference's expression.  Same arithmetic on host and device (this
// header); the host instantiation is what the CPU tests compare with the oracle's dense solve.
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

struct Sym2 { double a, b, d; };      // [[a, b], [b, d]]

struct Var1D {
    int M;
    double amp;
    const double* u;       // [M] sorted scaled training inputs
    const double* Pf;      // [3 M] filtered covariance given points <= k
    const double* Pb;      // [3 M] filtered covariance given points >= k, mirrored coordinates (f, -f')
};

// Pinf + A(d) (P - Pinf) A(d)^T
GPF_HD inline Sym2 var1d_predict(const Sym2& P, double d, double amp) {
    const double c = 1.7320508075688772, c2 = 3.0;
    const double e = exp(-c * d);
    const double a00 = e * (1.0 + c * d), a01 = e * d, a10 = -e * c2 * d, a11 = e * (1.0 - c * d);
    const double d00 = P.a - amp, d01 = P.b, d11 = P.d - c2 * amp;
    const double t00 = a00 * d00 + a01 * d01, t01 = a00 * d01 + a01 * d11;
    const double t10 = a10 * d00 + a11 * d01, t11 = a10 * d01 + a11 * d11;
    Sym2 R;
    R.a = amp + (t00 * a00 + t01 * a01);
    R.b = t00 * a10 + t01 * a11;
    R.d = c2 * amp + (t10 * a10 + t11 * a11);
    return R;
}

// posterior variance of f at the scaled position uq (without the conditioned GP's jitter)
GPF_HD inline double var1d_eval(const Var1D& T, double uq) {
    const double c2 = 3.0;
    int lo = 0, hi = T.M;                               // k = #{j : u_j <= uq}
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (T.u[mid] <= uq) lo = mid + 1; else hi = mid;
    }
    const int k = lo;
    Sym2 pf, pb;
    if (k > 0) {
        const Sym2 P = {T.Pf[3 * (k - 1)], T.Pf[3 * (k - 1) + 1], T.Pf[3 * (k - 1) + 2]};
        pf = var1d_predict(P, uq - T.u[k - 1], T.amp);
    }
    if (k < T.M) {
        const Sym2 P = {T.Pb[3 * k], T.Pb[3 * k + 1], T.Pb[3 * k + 2]};
        pb = var1d_predict(P, T.u[k] - uq, T.amp);
        pb.b = -pb.b;                                   // back from the mirrored coordinates
    }
    if (k == 0) return pb.a;
    if (k == T.M) return pf.a;
    const double df = pf.a * pf.d - pf.b * pf.b, db = pb.a * pb.d - pb.b * pb.b;
    const double i00 = pf.d / df + pb.d / db - 1.0 / T.amp;
    const double i01 = -pf.b / df - pb.b / db;
    const double i11 = pf.a / df + pb.a / db - 1.0 / (c2 * T.amp);
    return i11 / (i00 * i11 - i01 * i01);
}

// host-side tables: one Kalman sweep from either end of the sorted training points
struct Var1DHost {
    std::vector<double> u, Pf, Pb;
    double amp = 1.0;
    static void sweep(const std::vector<double>& t, double amp, double sig2, std::vector<double>& out) {
        const int M = (int)t.size();
        out.resize((size_t)3 * M);
        Sym2 P = {amp, 0.0, 3.0 * amp};
        for (int k = 0; k < M; ++k) {
            if (k > 0) P = var1d_predict(P, t[k] - t[k - 1], amp);
            const double S = P.a + sig2, r = sig2 / S;      // update with one observation of f, noise sig2
            Sym2 N;
            N.a = P.a * r;
            N.b = P.b * r;
            N.d = P.d - P.b * P.b / S;
            P = N;
            out[3 * k] = P.a; out[3 * k + 1] = P.b; out[3 * k + 2] = P.d;
        }
    }
    void build(const double* u_in, int M, double amp_, double yerr) {
        amp = amp_;
        u.assign(u_in, u_in + M);
        std::sort(u.begin(), u.end());
        sweep(u, amp, yerr * yerr, Pf);
        std::vector<double> t(M), pb;
        for (int k = 0; k < M; ++k) t[k] = -u[M - 1 - k];    // mirrored coordinates: the same filter from the right
        sweep(t, amp, yerr * yerr, pb);
        Pb.resize((size_t)3 * M);
        for (int k = 0; k < M; ++k)
            for (int e = 0; e < 3; ++e) Pb[3 * k + e] = pb[3 * (M - 1 - k) + e];
    }
    Var1D view() const {
        Var1D T;
        T.M = (int)u.size(); T.amp = amp; T.u = u.data(); T.Pf = Pf.data(); T.Pb = Pb.data();
        return T;
    }
};

}  // namespace gpf
