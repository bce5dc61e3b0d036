// Can a K-major SWIZZLE_128B UMMA A-operand start at an arbitrary ROW of a TMA-written slab (start address = slab + r*128 B,
// not 1024-byte aligned)?  Needed to keep the 128+6 activation rows of a k7 convolution tile in shared memory once and let
// the 7 taps read it at row offsets 0..6.  Tests descriptor base_offset = 0 and = r & 7 (bits 49-51) against the exact result.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I ace-step-1.5-ggml_b200/csrc scripts/probes/umma_rowshift_probe.cu -o /tmp/u -lcuda
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <vector>
#include "ptx.cuh"
using namespace ptx;

constexpr int SLAB_ROWS = 160;
__global__ void __launch_bounds__(128) probe(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, int shift, int bo, float* out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t full, done;
    __shared__ uint32_t slot;
    uint8_t* sA = smem;                      // 160 rows x 128 B
    uint8_t* sB = smem + 24576;              // 128 rows x 128 B (1024-aligned)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { mbar_init(&full, 1); mbar_init(&done, 1); fence_mbar_init(); }
    if (warp == 0) tmem_alloc<128>(&slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = slot;
    if (threadIdx.x == 0) {
        mbar_expect_tx(&full, SLAB_ROWS * 128 + 128 * 128);
        tma_load_3d(sA, &tmA, &full, 0, 0, 0);
        tma_load_3d(sB, &tmB, &full, 0, 0, 0);
        mbar_wait(&full, 0);
        tc_fence_after();
        constexpr uint32_t idesc = umma_idesc_f16(128, 128);
        const uint64_t da0 = umma_desc_sw128(smem_u32(sA) + shift * 128) | ((uint64_t)(bo & 7) << 49);
        const uint64_t db0 = umma_desc_sw128(smem_u32(sB));
        for (int k = 0; k < 4; ++k) umma_f16(tb, da0 + 2 * k, db0 + 2 * k, idesc, k ? 1u : 0u);
        umma_commit(&done);
        mbar_wait(&done, 0);
    }
    __syncthreads();
    tc_fence_after();
    for (int c = 0; c < 4; ++c) {
        uint32_t r[32];
        tmem_ld32(tb + ((uint32_t)(warp * 32) << 16) + c * 32, r);
        tmem_ld_wait();
        for (int i = 0; i < 32; ++i) out[(warp * 32 + lane) * 128 + c * 32 + i] = __uint_as_float(r[i]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) { tc_fence_after(); tmem_dealloc<128>(tb); }
}

typedef CUresult (*PFN_enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                            const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static bool make(PFN_enc enc, CUtensorMap* tm, void* base, int rows, int box_rows) {
    cuuint64_t dims[3] = {64, (cuuint64_t)rows, 1};
    cuuint64_t str[2] = {128, (cuuint64_t)rows * 128};
    cuuint32_t box[3] = {64, (cuuint32_t)box_rows, 1}, es[3] = {1, 1, 1};
    return enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, base, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
               CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

int main() {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    PFN_enc enc = (PFN_enc)fn;
    std::vector<__half> hA(SLAB_ROWS * 64), hB(128 * 64);
    for (int r = 0; r < SLAB_ROWS; ++r) for (int c = 0; c < 64; ++c) hA[r * 64 + c] = __float2half((float)((r * 7 + c * 3) % 61) - 30.f);
    for (int n = 0; n < 128; ++n) for (int k = 0; k < 64; ++k) hB[n * 64 + k] = __float2half(k == (n % 64) ? (n < 64 ? 1.f : 2.f) : 0.f);
    __half *dA, *dB;
    float* dO;
    cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dB, hB.size() * 2); cudaMalloc(&dO, 128 * 128 * 4);
    cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
    CUtensorMap tmA, tmB;
    if (!make(enc, &tmA, dA, SLAB_ROWS, SLAB_ROWS) || !make(enc, &tmB, dB, 128, 128)) { printf("encode failed\n"); return 1; }
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 49152);
    std::vector<float> o(128 * 128);
    for (int shift : {0, 1, 2, 3, 5, 6, 8, 9, 18, 27}) {
        for (int mode = 0; mode < 2; ++mode) {
            const int bo = mode ? (shift & 7) : 0;
            cudaMemset(dO, 0, o.size() * 4);
            probe<<<1, 128, 49152>>>(tmA, tmB, shift, bo, dO);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("shift %d bo %d: kernel failed: %s\n", shift, bo, cudaGetErrorString(e)); return 1; }
            cudaMemcpy(o.data(), dO, o.size() * 4, cudaMemcpyDeviceToHost);
            int bad = 0;
            for (int m = 0; m < 128; ++m) for (int n = 0; n < 128; ++n) {
                const float want = (float)(((m + shift) * 7 + (n % 64) * 3) % 61 - 30) * (n < 64 ? 1.f : 2.f);
                if (o[m * 128 + n] != want) ++bad;
            }
            printf("row shift %2d, base_offset %d: %s (%d wrong of 16384)\n", shift, bo, bad ? "WRONG" : "exact", bad);
        }
    }
    return 0;
}
