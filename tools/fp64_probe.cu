// FP64 pipe probe for B200: DFMA throughput vs resident warps and ILP, and whether integer
// instructions issue under a busy FP64 pipe.   nvcc -arch=sm_100a -O3 -o fp64_probe fp64_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
template <int ILP, int NINT>
__global__ void k(double* out, int iters, double a, double b) {
    double x[ILP];
    int z[4] = {threadIdx.x, 1, 2, 3};
#pragma unroll
    for (int i = 0; i < ILP; ++i) x[i] = a + i + threadIdx.x;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r) {
#pragma unroll
            for (int i = 0; i < ILP; ++i) x[i] = __fma_rn(x[i], a, b);
#pragma unroll
            for (int i = 0; i < NINT; ++i) z[i & 3] = (z[i & 3] ^ (z[(i + 1) & 3] + it)) * 3;
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + z[0] + z[1] + z[2] + z[3];
}
template <int ILP, int NINT>
void run(int warps_per_sm, double* d) {
    int dev; cudaGetDevice(&dev); int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, dev);
    const int iters = 20000;
    dim3 grid(sms), block(32 * warps_per_sm);
    k<ILP, NINT><<<grid, block>>>(d, 10, 1.0000001, 1e-9);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<ILP, NINT><<<grid, block>>>(d, iters, 1.0000001, 1e-9);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double cycles = ms * 1e-3 * clk * 1e3;
    double dfma = (double)iters * 8 * ILP * warps_per_sm;           // warp-instructions per SM
    double ints = (double)iters * 8 * NINT * 3 * warps_per_sm;      // ~3 integer instructions per z update
    printf("warps/SM %2d ILP %d int/fp %d : %.3f DFMA warp-inst/clk/SM  (%.3f int/clk/SM)  %.2f ms\n", warps_per_sm, ILP, NINT,
           dfma / cycles, ints / cycles, ms);
}
int main() {
    double* d; cudaMalloc(&d, 148 * 1024 * 8 * 4);
    for (int w : {4, 8, 16, 32}) { run<1, 0>(w, d); run<2, 0>(w, d); run<4, 0>(w, d); run<8, 0>(w, d); }
    for (int w : {8, 16, 32}) { run<4, 1>(w, d); run<4, 2>(w, d); run<4, 4>(w, d); }
    return 0;
}
