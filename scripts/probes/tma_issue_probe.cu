// How fast can ONE thread run the producer handshake of the tap-GEMM?  Per iteration: wait(empty-like barrier) + expect_tx +
// NOPS TMA operations (box loads or 1-D bulk copies of `bytes` each) on a 4-deep ring.  Reports cycles per iteration.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I ace-step-1.5-ggml_b200/csrc scripts/probes/tma_issue_probe.cu -o /tmp/p -lcuda
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <vector>
#include "ptx.cuh"
using namespace ptx;

constexpr int ST = 4;

template <int NOPS, bool BOX>
__global__ void __launch_bounds__(64) probe(const __grid_constant__ CUtensorMap tm, const __half* base, int rows, int iters, long long* cycles) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t full[ST];
    const uint32_t bytes = (uint32_t)rows * 128u;
    if (threadIdx.x == 0) {
        for (int s = 0; s < ST; ++s) mbar_init(&full[s], 1);
        fence_mbar_init();
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    const int row0 = blockIdx.x * 1024;
    long long t0 = clock64();
    for (int i = 0; i < iters + ST; ++i) {
        const int s = i & (ST - 1);
        if (i >= ST) mbar_wait(&full[s], ((i >> 2) - 1) & 1);
        if (i < iters) {
            mbar_expect_tx(&full[s], bytes * NOPS);
            const int r = row0 + ((i * 8) & 511);
#pragma unroll
            for (int o = 0; o < NOPS; ++o) {
                uint8_t* dst = smem + s * 49152 + o * (49152 / NOPS);
                if (BOX) tma_load_3d(dst, &tm, &full[s], (o & 1) * 64, r, 0);
                else bulk_load(dst, base + (size_t)(r + o * 128) * 128, bytes, &full[s]);
            }
        }
    }
    cycles[blockIdx.x] = clock64() - t0;
}

typedef CUresult (*PFN_enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                            const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <int NOPS, bool BOX>
void run(const CUtensorMap& tm, const __half* d, int rows, int grid, long long* cyc) {
    const int iters = 4000;
    cudaFuncSetAttribute(probe<NOPS, BOX>, cudaFuncAttributeMaxDynamicSharedMemorySize, ST * 49152);
    for (int rep = 0; rep < 2; ++rep) probe<NOPS, BOX><<<grid, 64, ST * 49152>>>(tm, d, rows, iters, cyc);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(e)); exit(1); }
    std::vector<long long> h(grid);
    cudaMemcpy(h.data(), cyc, grid * 8, cudaMemcpyDeviceToHost);
    double s = 0;
    for (auto v : h) s += (double)v;
    const double per = s / grid / iters;
    printf("%s ops/iter %d x %5d B grid %3d: %7.1f cycles/iter  %6.1f B/clk/SM\n", BOX ? "box " : "bulk", NOPS, rows * 128, grid, per, NOPS * rows * 128 / per);
}

int main() {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    PFN_enc enc = (PFN_enc)fn;
    const int C = 128, nsm = 148;
    const size_t n = (size_t)nsm * 1024 * C;
    __half* d;
    cudaMalloc(&d, n * 2);
    cudaMemset(d, 0, n * 2);
    long long* cyc;
    cudaMalloc(&cyc, nsm * 8);
    for (int rows : {8, 64, 128}) {
        CUtensorMap tm;
        cuuint64_t dims[3] = {(cuuint64_t)C, (cuuint64_t)nsm * 1024, 1};
        cuuint64_t str[2] = {(cuuint64_t)C * 2, (cuuint64_t)n * 2};
        cuuint32_t box[3] = {64, (cuuint32_t)rows, 1}, es[3] = {1, 1, 1};
        if (enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, d, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) return 1;
        for (int grid : {1, 148}) {
            run<1, true>(tm, d, rows, grid, cyc);
            run<2, true>(tm, d, rows, grid, cyc);
            run<3, true>(tm, d, rows, grid, cyc);
            run<1, false>(tm, d, rows, grid, cyc);
            run<2, false>(tm, d, rows, grid, cyc);
            run<3, false>(tm, d, rows, grid, cyc);
        }
    }
    return 0;
}
