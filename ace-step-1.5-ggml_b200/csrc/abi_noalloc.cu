// ace_b200_attention_noalloc: the attention kernel of the DiT path (attention_tc.cu; reference semantics:
// acestep_dit_model.cpp:1212-1250, ggml soft_max_ext over Q.K^T with an optional key mask / sliding window) for callers that
// must not allocate device memory — code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc.
// It has no counterpart in the reference.
//
// The existing entry point (ace_b200_debug_attention) cudaMallocs its fp16 operand copies per call.  Here that scratch is a
// file-scope `static __device__` array of NOALLOC_SCRATCH_BYTES (16 MiB): every GPU that loads the library holds it until exit.
//   * This entry point calls no cudaMalloc / cudaFree / allocating library routine; sizes beyond the array are rejected with
//     ACE_GGML_ERR_INVALID_ARG and a message that states the need and the limit.
//   * In HOST code the array's name is a host-side stand-in, not a device address.  The device address is fetched once per
//     device, with that device current, by cudaGetSymbolAddress (return code checked) and only that pointer reaches kernels.
//   * Every call on a GPU shares the array, and it keeps the last call's contents.  Each call overwrites every byte it later
//     reads (operands, mask) and clears the output region first, so nothing leaks from call to call; calls that overlap on one
//     GPU (two contexts / streams / threads on the same device) RACE on it — serialise them.
#include "abi_internal.h"

#include <mutex>

using namespace ace_abi;

namespace {

constexpr size_t NOALLOC_SCRATCH_BYTES = ACE_B200_NOALLOC_SCRATCH_BYTES;  // include/acestep_b200.h: 16 MiB
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

std::mutex g_scratch_mu;
unsigned char* g_scratch_ptr[ace::ACE_MAX_DEVICES] = {};

// device address of the array on `dev` (which must be the current device)
cudaError_t scratch_address(int dev, unsigned char** out) {
    if (dev < 0 || dev >= ace::ACE_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(g_scratch_mu);
    if (!g_scratch_ptr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        g_scratch_ptr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_scratch_ptr[dev];
    return cudaSuccess;
}

__global__ void f16_to_f32_rows_kernel(const __half* __restrict__ src, float* __restrict__ dst, long n) {
    const long i = ((long)blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (i + 1 < n) {
        const float2 v = __half22float2(*reinterpret_cast<const __half2*>(src + i));
        *reinterpret_cast<float2*>(dst + i) = v;
    } else if (i < n) {
        dst[i] = __half2float(src[i]);
    }
}

inline size_t up256(size_t b) { return (b + 255) & ~(size_t)255; }

}  // namespace

extern "C" {

ace_ggml_status ace_b200_attention_noalloc(ace_ggml_context* ctx, const float* Q, const float* K, const float* V, const int32_t* key_mask,
                                           int32_t B, int32_t q_len, int32_t k_len, int32_t n_heads, int32_t n_kv, int32_t window, float* out,
                                           size_t out_size) {
    if (!ctx || !Q || !K || !V || !out || B <= 0 || q_len <= 0 || k_len <= 0 || n_heads <= 0 || n_kv <= 0 || n_heads % n_kv) return ACE_GGML_ERR_INVALID_ARG;
    DeviceGuard guard(ctx->device);
    const size_t nq = (size_t)B * q_len * n_heads * 128, nk = (size_t)B * k_len * n_kv * 128;
    if (out_size < nq * 4) return set_error(ctx, ACE_GGML_ERR_INVALID_ARG, "output buffer too small");
    // carve: [fp32 staging | Q16 | K16 | V16 | O16 | mask]
    const size_t b32 = up256(std::max(nq, nk) * 4), bq = up256(nq * 2), bk = up256(nk * 2), bm = key_mask ? up256((size_t)B * k_len * 4) : 0;
    const size_t need = b32 + 2 * bq + 2 * bk + bm;
    if (need > NOALLOC_SCRATCH_BYTES)
        return set_error(ctx, ACE_GGML_ERR_INVALID_ARG, "ace_b200_attention_noalloc: needs " + std::to_string(need) + " bytes of scratch, the static array holds " +
                                                            std::to_string(NOALLOC_SCRATCH_BYTES) + " (use the allocating entry point for this size)");
    unsigned char* base = nullptr;
    const cudaError_t ea = scratch_address(ctx->device, &base);
    if (ea != cudaSuccess) return cuda_fail(ctx, "ace_b200_attention_noalloc: cudaGetSymbolAddress failed", ea);
    float* d32 = reinterpret_cast<float*>(base);
    __half* dQ = reinterpret_cast<__half*>(base + b32);
    __half* dK = reinterpret_cast<__half*>(base + b32 + bq);
    __half* dV = reinterpret_cast<__half*>(base + b32 + bq + bk);
    __half* dO = reinterpret_cast<__half*>(base + b32 + bq + 2 * bk);
    int32_t* dM = key_mask ? reinterpret_cast<int32_t*>(base + b32 + 2 * bq + 2 * bk) : nullptr;

    cudaStream_t st = ctx->stream;
    cudaMemsetAsync(dO, 0, nq * 2, st);  // the array keeps the last call's output: never let it show through
    cudaMemcpyAsync(d32, Q, nq * 4, cudaMemcpyDefault, st);
    launch_f32_to_f16(d32, dQ, (long)nq, st);
    cudaMemcpyAsync(d32, K, nk * 4, cudaMemcpyDefault, st);
    launch_f32_to_f16(d32, dK, (long)nk, st);
    cudaMemcpyAsync(d32, V, nk * 4, cudaMemcpyDefault, st);
    launch_f32_to_f16(d32, dV, (long)nk, st);
    if (key_mask) cudaMemcpyAsync(dM, key_mask, (size_t)B * k_len * 4, cudaMemcpyDefault, st);
    cudaEventRecord(ctx->ev0, st);
    launch_attention(dQ, (long)n_heads * 128, dK, (long)n_kv * 128, dV, (long)n_kv * 128, dO, (long)n_heads * 128, dM, B, q_len, k_len,
                     n_heads, n_kv, window, st);
    cudaEventRecord(ctx->ev1, st);
    note_launch();
    f16_to_f32_rows_kernel<<<(unsigned)((nq / 2 + 256) / 256), 256, 0, st>>>(dO, d32, (long)nq);
    cudaMemcpyAsync(out, d32, nq * 4, cudaMemcpyDefault, st);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return cuda_fail(ctx, "ace_b200_attention_noalloc failed", e);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->last_device_ms = ms;
    return ACE_GGML_OK;
}

}  // extern "C"
