// Host-side GP model object behind the opaque gpf_gp handle.
#pragma once
#include <cuda_runtime.h>
#include <vector>
#include "gpf_types.h"
#include "gp_mean1d.h"
#include "gp_var1d.h"

namespace gpf {

struct GpKern {          // amp * Matern32(|| s . (x - x') ||)
    int D;
    double amp;
    double s[8];         // exp(-log_scale)
};

struct FvSolver;

struct GpModel {
    int Mmax, D, nout, M, Mp, nslots;
    cudaStream_t side = nullptr;          // second stream of the factorisation (trailing update beside the next panel)
    cudaEvent_t ev_strip = nullptr, ev_rest = nullptr;
    bool fast_arith = true;   // gpf_gp_set_arith: contracted exp / sqrt in the mean kernel (1e-12; the GP tolerance is 1e-9)
    GpKern kp;
    // carved from the caller's workspace
    double* Xs;          // (Mp, D) pre-scaled training inputs
    double* L;           // (Mp, Mp) row-major lower Cholesky factor (identity in the padding)
    double* Dinv;        // (Mp/64, 64, 64) inverses of L's diagonal blocks
    double* alpha;       // (2, Mp) K^-1 Y, one row per output
    double* ycopy;       // (2, Mp) the right-hand sides Y (training loss)
    double* scratch;     // (nslots, Mp, 128) v panels of the variance kernel
    double* gpart;       // partial maxima of the mean kernel
    double* vpart;       // per-CTA variance maxima
    long long* ipart;    // per-CTA arg-maxima
    double* scal;        // result scalars
    // one-input, one-output models (the pressure GP on the density): O(log M) mean (gp_mean1d.h)
    double* t1d;         // 5 Mp doubles: u, A, B, C, D
    int* t1d_i;          // Mp + Mp + 1 ints: blk, bstart
    std::vector<double> u_host;      // scaled training inputs of the last factorisation (D == 1)
    bool t1d_valid = false;          // tables match the last factorisation
    Mean1D t1d_view;
    int build_mean1d(cudaStream_t s);
    // O(log M) variance of one-input models (gp_var1d.h); GPF_GP_VAR1D=0 restores the dense solve
    double* v1d;         // 7 Mp doubles: u, Pf (3), Pb (3)
    double yerr_last = 0.0;
    bool v1d_valid = false;
    Var1D v1d_view;
    int build_var1d(cudaStream_t s);

    static size_t workspace_bytes(int Mmax, int D, int nout, int nslots);
    static int var_slots(int sms);           // resident CTAs of the variance kernel = v panels in the workspace
    int init(int Mmax, int D, int nout, void* ws, size_t bytes);
    int factor(const double* Xtrain, int M, double log_amp, const double* log_scale, double yerr, const double* Y,
               cudaStream_t s);
    int get_factor(double* L_out, double* alpha_out, cudaStream_t s);
    static size_t train_workspace_bytes(int Mmax);
    int nlml(void* work, size_t bytes, double* loss, double* grad, cudaStream_t s);
    int mean(const double* Xtest, long long N, double* mean_out, double* dmax_host, cudaStream_t s);
    int mean_var(const double* Xtest, long long N, const double* noise, double noise_scalar, double* mean_out,
                 double* var_out, double* vmax_host, long long* imax_host, cudaStream_t s);
};

int launch_gp_features(FvSolver& S, const double* x_scale7, const int* dims, int D, double* X, cudaStream_t s, int rel = 0);

}  // namespace gpf
