// L2 -> shared-memory delivery rate of TMA on B200, per SM, for the access shapes the tap-GEMM uses.
//   mode 0: 3-D tensor-map boxes 64 x ROWS fp16 (128 B x ROWS rows, SWIZZLE_128B), every CTA its own L2-resident region
//   mode 1: same boxes, every CTA the SAME region
//   mode 2: 1-D bulk copies of the same byte count, own region        mode 3: 1-D bulk, same region
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I ace-step-1.5-ggml_b200/csrc scripts/probes/tma_probe.cu -o gpurun_out/tma_probe -lcuda
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "ptx.cuh"
using namespace ptx;

constexpr int MAXST = 12;

__global__ void __launch_bounds__(64) probe(const __grid_constant__ CUtensorMap tm, const __half* base, int mode, int rows, int pitch_elems,
                                            int rows_per_cta, int iters, long long* cycles, int STAGES, int slot) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t full[MAXST];
    const uint32_t bytes = (uint32_t)rows * 128u;
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) mbar_init(&full[s], 1);
        fence_mbar_init();
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    const int cta = (mode & 1) ? 0 : blockIdx.x;
    const int row0 = cta * rows_per_cta;
    long long t0 = clock64();
    for (int i = 0; i < iters + STAGES; ++i) {
        const int s = i % STAGES;
        if (i >= STAGES) mbar_wait(&full[s], ((i / STAGES) - 1) & 1);
        if (i < iters) {
            mbar_expect_tx(&full[s], bytes);
            // walk the CTA's region: tap-like shifts of a few rows and the two 64-column halves
            const int r = row0 + (i * 3) % (rows_per_cta - rows);
            const int c = (i & 1) * 64;
            if (mode < 2) tma_load_3d(smem + s * slot, &tm, &full[s], c, r, 0);
            else bulk_load(smem + s * slot, base + (size_t)(row0 + ((i * 128) % (rows_per_cta - rows))) * pitch_elems, bytes, &full[s]);
        }
    }
    cycles[blockIdx.x] = clock64() - t0;
}

typedef CUresult (*PFN_enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                            const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    PFN_enc enc = (PFN_enc)fn;
    const int C = 128, rows_per_cta = 1024, nsm = 148;
    const size_t n = (size_t)nsm * rows_per_cta * C;
    __half* d;
    cudaMalloc(&d, n * 2);
    cudaMemset(d, 0, n * 2);
    long long* cyc;
    cudaMalloc(&cyc, nsm * 8);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 6 * 32768);
    for (int rows : {32, 128, 256}) {
        CUtensorMap tm;
        cuuint64_t dims[3] = {(cuuint64_t)C, (cuuint64_t)nsm * rows_per_cta, 1};
        cuuint64_t str[2] = {(cuuint64_t)C * 2, (cuuint64_t)n * 2};
        cuuint32_t box[3] = {64, (cuuint32_t)rows, 1}, es[3] = {1, 1, 1};
        CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, d, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
        for (int grid : {148, 1}) for (int STAGES : {2, 4, 6, 12}) {
            const int slot = rows * 128;
            if (STAGES * slot > 6 * 32768) continue;
            for (int mode = 0; mode < 4; ++mode) {
                const int iters = 2000;
                for (int rep = 0; rep < 2; ++rep) probe<<<grid, 64, 6 * 32768>>>(tm, d, mode, rows, C, rows_per_cta, iters, cyc, STAGES, slot);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(e)); return 1; }
                std::vector<long long> h(grid);
                cudaMemcpy(h.data(), cyc, grid * 8, cudaMemcpyDeviceToHost);
                double s = 0;
                for (auto v : h) s += (double)v;
                s /= grid;
                const double per = s / iters;
                printf("stages %2d rows %3d (%5d B) grid %3d mode %d (%s, %s): %7.1f cycles/load  %6.1f B/clk/SM  %7.0f B/clk chip\n", STAGES, rows, rows * 128, grid, mode,
                       mode < 2 ? "box " : "bulk", (mode & 1) ? "same" : "own ", per, rows * 128 / per, rows * 128 / per * grid);
            }
        }
    }
    return 0;
}
