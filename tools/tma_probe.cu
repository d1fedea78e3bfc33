// Probe: which TMA box/tensor shapes load correctly for FP64 planes on this GPU/driver.
// usage: tma_probe <rank> <box0> <box1> <box2> <c0> <c1> <c2> [u64]
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__global__ void probe(const __grid_constant__ CUtensorMap m, int rank, int c0, int c1, int c2, unsigned bytes, double* out, int n) {
    extern __shared__ __align__(128) unsigned char sm[];
    __shared__ __align__(8) unsigned long long bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(bytes) : "memory");
        if (rank == 3)
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                         ::"r"(smem_u32(sm)), "l"((unsigned long long)&m), "r"(smem_u32(&bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
        else
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                         ::"r"(smem_u32(sm)), "l"((unsigned long long)&m), "r"(smem_u32(&bar)), "r"(c0), "r"(c1) : "memory");
    }
    asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra D;\nbra W;\nD:\n}\n" ::"r"(smem_u32(&bar)) : "memory");
    for (int i = threadIdx.x; i < n; i += blockDim.x) out[i] = ((double*)sm)[i];
}
int main(int argc, char** argv) {
    int rank = atoi(argv[1]); cuuint32_t box[3] = {(cuuint32_t)atoi(argv[2]), (cuuint32_t)atoi(argv[3]), (cuuint32_t)atoi(argv[4])};
    int c0 = atoi(argv[5]), c1 = atoi(argv[6]), c2 = atoi(argv[7]);
    bool u64 = argc > 8;
    const int pitch = 80, rows = 102; size_t plane = (size_t)pitch * rows;
    std::vector<double> h(3 * plane);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (double)i;
    double* d; cudaMalloc(&d, h.size() * 8); cudaMemcpy(d, h.data(), h.size() * 8, cudaMemcpyHostToDevice);
    void* fp = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q);
    EncodeTiledFn enc = (EncodeTiledFn)fp;
    CUtensorMap m; cuuint64_t dims[3] = {pitch, rows, 3}; cuuint64_t strides[2] = {pitch * 8, plane * 8}; cuuint32_t es[3] = {1, 1, 1};
    CUresult r = enc(&m, u64 ? CU_TENSOR_MAP_DATA_TYPE_UINT64 : CU_TENSOR_MAP_DATA_TYPE_FLOAT64, rank, d, dims, strides, box, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode rc=%d ", (int)r);
    int n = box[0] * box[1] * (rank == 3 ? box[2] : 1);
    double* o; cudaMalloc(&o, n * 8);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 100000);
    probe<<<1, 128, n * 8 + 128>>>(m, rank, c0, c1, c2, n * 8, o, n);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel: %s ", cudaGetErrorString(e));
    if (e == cudaSuccess) {
        std::vector<double> ho(n); cudaMemcpy(ho.data(), o, n * 8, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int z = 0; z < (rank == 3 ? (int)box[2] : 1); ++z) for (int y = 0; y < (int)box[1]; ++y) for (int x = 0; x < (int)box[0]; ++x) {
            int gx = c0 + x, gy = c1 + y, gz = c2 + z;
            double want = (gx < 0 || gx >= pitch || gy < 0 || gy >= rows || gz >= 3) ? 0.0 : (double)(gz * plane + gy * pitch + gx);
            if (ho[(z * box[1] + y) * box[0] + x] != want) bad++;
        }
        printf("mismatches=%d", bad);
    }
    printf("\n");
    return 0;
}
