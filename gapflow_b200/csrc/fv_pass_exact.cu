// Pass kernels, bit-exact arithmetic.  Compile with -fmad=false (gapflow_b200/build.py).
#define GPF_ARITH_NS exact
#define GPF_FAST_ARITH 0
#include "fv_pass_kernel.cuh"

namespace gpf {
int launch_pass_exact(const PassArgs& A, bool corr, bool ls, bool dh, bool xf, bool vf, int n, cudaStream_t s) {
    return corr ? exact::launch_pass_dispatch<true>(A, ls, dh, xf, vf, n, s) : exact::launch_pass_dispatch<false>(A, ls, dh, xf, vf, n, s);
}
}  // namespace gpf
