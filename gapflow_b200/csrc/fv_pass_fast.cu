// Pass kernels, contracted arithmetic.  Compile with -fmad=true (gapflow_b200/build.py).
#define GPF_ARITH_NS fastm
#define GPF_FAST_ARITH 1
#include "fv_pass_kernel.cuh"

namespace gpf {
int launch_pass_fast(const PassArgs& A, bool corr, bool ls, bool dh, bool xf, bool vf, int n, cudaStream_t s) {
    return corr ? fastm::launch_pass_dispatch<true>(A, ls, dh, xf, vf, n, s) : fastm::launch_pass_dispatch<false>(A, ls, dh, xf, vf, n, s);
}
}  // namespace gpf
