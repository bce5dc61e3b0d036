// What slows a tcgen05.mma stream (M=128, N=128/256, K=16, operands in shared memory) when the rest of the SM is busy?
// Warp 1 lane 0 issues `iters` MMAs walking 4 operand stages; the other warps run one kind of background work until it is done:
//   noise 0: none   1: FFMA loops (8 warps)   2: shared-memory st/ld streams (8 warps)   3: tcgen05.ld of the accumulator (4 warps)
//   4: one thread streaming 1-D bulk copies (L2 -> smem, 2 x 16 KB per 300 cycles) into a separate ring   5: 1 + 2 + 3 + 4
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I ace-step-1.5-ggml_b200/csrc scripts/probes/mma_contention_probe.cu -o /tmp/c
#include <cuda_runtime.h>
#include <cstdio>
#include <vector>
#include "ptx.cuh"
using namespace ptx;

constexpr int OPER = 3 * 49152;         // MMA operand stages
constexpr int RING = 2 * 16384;         // TMA noise ring
constexpr int STG = 8 * 4096;           // LSU noise buffers
template <int N>
__global__ void __launch_bounds__(320) probe(int noise, int iters, const uint8_t* gsrc, long long* cycles, float* sink, long long* ldcount, int random_data) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar, nbar[2], ready, cyc_bar;
    __shared__ uint32_t slot;
    __shared__ volatile int done;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool handshake = noise >= 10;
    if (handshake) noise -= 10;
    for (int i = threadIdx.x; i < OPER / 4; i += blockDim.x) {  // operands: all 1.0, or random fp16 in (-2, 2) (data toggling = power)
        uint32_t h = (uint32_t)i * 2654435761u + blockIdx.x * 40503u;
        h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
        reinterpret_cast<uint32_t*>(smem)[i] = random_data ? ((h & 0x83ff83ffu) | 0x3c003c00u) : 0x3c003c00u;
    }
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_init(&nbar[0], 1); mbar_init(&nbar[1], 1); mbar_init(&ready, 1); mbar_init(&cyc_bar, 1); fence_mbar_init(); mbar_arrive(&ready); done = 0; }
    if (warp == 0) tmem_alloc<512>(&slot);
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = slot;
    if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_f16(128, N);
            const uint64_t da0 = umma_desc_sw128(smem_u32(smem)), db0 = umma_desc_sw128(smem_u32(smem) + 16384);
            long long t0 = clock64();
            uint32_t s = 0;
            if (iters < 0) while (clock64() - t0 < 2000000) {}
            for (int i = 0; i < iters; i += 4) {
                const uint64_t da = da0 + (uint64_t)(s * (49152 >> 4)), db = db0 + (uint64_t)(s * (49152 >> 4));
                if (handshake) {   // what the GEMM main loop does around every k-block
                    mbar_wait(&ready, 0);
                    tc_fence_after();
                }
#pragma unroll
                for (int k = 0; k < 4; ++k) umma_f16(tb, da + 2 * k, db + 2 * k, idesc, (i | k) ? 1u : 0u);
                if (handshake) umma_commit(&cyc_bar);
                s = (s == 2) ? 0 : s + 1;
            }
            umma_commit(&bar);
            const long long t1 = clock64();
            mbar_wait(&bar, 0);
            const long long t2 = clock64();
            cycles[blockIdx.x * 2] = t1 - t0;
            cycles[blockIdx.x * 2 + 1] = t2 - t0;
            done = 1;
        }
    } else if (warp == 0) {
        if (lane == 0 && (noise == 4 || noise == 5)) {
            uint32_t i = 0;
            while (!done) {
                const uint32_t b = i & 1;
                if (i >= 2) mbar_wait(&nbar[b], ((i >> 1) - 1) & 1);
                mbar_expect_tx(&nbar[b], 16384);
                bulk_load(smem + OPER + b * 16384, gsrc + (size_t)blockIdx.x * 65536 + (i & 3) * 16384, 16384, &nbar[b]);
                ++i;
            }
            for (uint32_t j = (i >= 2 ? i - 2 : 0); j < i; ++j) mbar_wait(&nbar[j & 1], (j >> 1) & 1);
        }
    } else {
        const int w = warp - 2;  // 0..7
        float acc = lane;
        if (noise == 1 || noise == 5) {
            float a = 1.0001f, b = 0.5f;
            while (!done) {
#pragma unroll
                for (int i = 0; i < 64; ++i) acc = fmaf(acc, a, b);
            }
        }
        if (noise == 2 || (noise == 5 && w >= 4)) {
            float4* buf = reinterpret_cast<float4*>(smem + OPER + RING + w * 4096);
            float4 v = make_float4(acc, 1.f, 2.f, 3.f);
            while (!done) {
#pragma unroll
                for (int i = 0; i < 8; ++i) buf[lane * 8 + (i ^ (lane & 7))] = v;
                __syncwarp();
#pragma unroll
                for (int i = 0; i < 8; ++i) { const float4 t = buf[(4 * i + (lane >> 3)) * 8 + ((lane & 7) ^ ((4 * i + (lane >> 3)) & 7))]; v.x += t.y; }
                __syncwarp();
            }
            acc += v.x;
        }
        if ((noise == 3 || noise == 5) && w < 4) {
            uint32_t r[32];
            long long n = 0;
            while (!done) {
                tmem_ld32(tb + ((uint32_t)(w * 32) << 16) + 256, r);   // the accumulator buffer the MMAs do NOT write
                tmem_ld_wait();
                acc += __uint_as_float(r[lane]);
                ++n;
            }
            if (lane == 0) ldcount[blockIdx.x * 4 + w] = n;
        }
        if (acc == 123.456f) sink[threadIdx.x] = acc;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(tb); }
}

template <int N>
void run(int grid, const uint8_t* g, long long* cyc, float* sink, long long* ldc, int rnd, int iters_arg) {
    const int smem_bytes = OPER + RING + STG;
    cudaFuncSetAttribute(probe<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    for (int noise : {0, 1, 2, 3, 4, 5, 10, 15}) {
        const int iters = iters_arg;
        if (rnd && noise != 0 && noise != 5) continue;
        if (noise >= 10 && rnd) continue;
        for (int rep = 0; rep < 2; ++rep) probe<N><<<grid, 320, smem_bytes>>>(noise, iters, g, cyc, sink, ldc, rnd);
        cudaError_t e = cudaGetLastError(); if (e == cudaSuccess) e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("failed: %s\n", cudaGetErrorString(e)); exit(1); }
        std::vector<long long> h(2 * grid);
        cudaMemcpy(h.data(), cyc, 2 * grid * 8, cudaMemcpyDeviceToHost);
        double a = 0, b = 0;
        for (int i = 0; i < grid; ++i) { a += h[2 * i]; b += h[2 * i + 1]; }
        printf("%s N=%3d grid %3d noise %d: issue %6.1f cycles/MMA, complete %6.1f cycles/MMA (tensor-pipe floor %d)\n", rnd ? "random operands, 2M MMAs" : "all-1.0 operands", N, grid, noise, a / grid / iters,
               b / grid / iters, N / 2);
        if (noise == 3) {  // how fast do the tcgen05.ld warps run next to the MMA stream, and with the tensor pipe idle?
            std::vector<long long> hc(4 * grid);
            cudaMemcpy(hc.data(), ldc, hc.size() * 8, cudaMemcpyDeviceToHost);
            double n = 0;
            for (auto v : hc) n += (double)v;
            const double busy = (b / grid) / (n / (4 * grid));
            probe<N><<<grid, 320, smem_bytes>>>(noise, -1, g, cyc, sink, ldc, rnd);
            cudaDeviceSynchronize();
            cudaMemcpy(hc.data(), ldc, hc.size() * 8, cudaMemcpyDeviceToHost);
            n = 0;
            for (auto v : hc) n += (double)v;
            printf("      tcgen05.ld 32x32b.x32 + wait, 4 warps: %.0f cycles per load next to the MMA stream, %.0f with the tensor pipe idle\n", busy,
                   2000000.0 / (n / (4 * grid)));
        }
    }
}

int main() {
    long long* cyc;
    cudaMalloc(&cyc, 148 * 16);
    float* sink;
    cudaMalloc(&sink, 4096);
    uint8_t* g;
    cudaMalloc(&g, 148 * 65536);
    cudaMemset(g, 0, 148 * 65536);
    long long* ldc;
    cudaMalloc(&ldc, 148 * 4 * 8);
    run<128>(148, g, cyc, sink, ldc, 0, 20000);
    run<256>(148, g, cyc, sink, ldc, 0, 20000);
    run<128>(148, g, cyc, sink, ldc, 1, 2000000);
    run<256>(148, g, cyc, sink, ldc, 1, 2000000);
    return 0;
}
