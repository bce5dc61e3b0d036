// Unit-test hooks of the C ABI (include/acestep_b200.h, "debug_*"): direct access to the tap-GEMM and attention
// kernels and to the weight packer, so tests/ can compare each against the oracle through the shared library.
#include "abi_internal.h"
#include "tensor_file.h"

#include <memory>

using namespace ace_abi;

namespace {
struct TmpAllocs {
    std::vector<void*> v;
    void* get(size_t bytes) {
        void* p = nullptr;
        if (cudaMalloc(&p, bytes + 256) != cudaSuccess) return nullptr;
        v.push_back(p);
        return p;
    }
    ~TmpAllocs() {
        for (void* p : v) cudaFree(p);
    }
};
}  // namespace

extern "C" {

ace_ggml_status ace_b200_debug_pack_roundtrip(const void* W, int32_t w_ggml_type, int32_t N, int32_t K, float* unpacked, float* ref) {
    if (!W || !unpacked || !ref || N <= 0 || K <= 0) return ACE_GGML_ERR_INVALID_ARG;
    std::vector<uint8_t> img;
    PackedWeight meta;
    std::string err;
    if (w_ggml_type == GGML_F32) {
        if (!pack_tiles_f32(static_cast<const float*>(W), N, 1, K, W_F16, img, meta, err)) return ACE_GGML_ERR;
        for (size_t i = 0; i < (size_t)N * K; ++i) ref[i] = f16_to_f32(f32_to_f16(static_cast<const float*>(W)[i]));
    } else {
        if (!pack_tiles_ggml(W, w_ggml_type, N, K, img, meta, err)) return ACE_GGML_ERR;
        const size_t rb = ggml_row_bytes(w_ggml_type, K);
        for (int r = 0; r < N; ++r)
            if (!dequant_row(w_ggml_type, static_cast<const uint8_t*>(W) + (size_t)r * rb, ref + (size_t)r * K, K)) return ACE_GGML_ERR;
    }
    std::vector<float> w;
    unpack_tiles_host(img, meta, w);
    std::memcpy(unpacked, w.data(), (size_t)N * K * 4);
    return ACE_GGML_OK;
}

ace_ggml_status ace_b200_debug_gemm(ace_ggml_context* ctx, const float* A, int32_t B, int32_t L, int32_t C, const void* W,
                                    int32_t w_ggml_type, int32_t N, int32_t taps, int32_t tap_stride, int32_t tap_offset, int32_t l_out,
                                    const float* bias, const float* resid, int32_t epilogue, int32_t simt, float* out, size_t out_size) {
    if (!ctx || !A || !W || !out || B <= 0 || L <= 0 || C <= 0 || N <= 0 || taps <= 0 || l_out <= 0) return ACE_GGML_ERR_INVALID_ARG;
    DeviceGuard guard(ctx->device);
    const bool swiglu = (epilogue & 2) != 0;
    const int n_out = swiglu ? N / 2 : N;
    const size_t out_elems = (size_t)B * l_out * n_out;
    if (out_size < out_elems * 4) return set_error(ctx, ACE_GGML_ERR_INVALID_ARG, "output buffer too small");
    std::vector<uint8_t> img;
    PackedWeight pw;
    std::string err;
    bool ok;
    if (w_ggml_type == GGML_F32) ok = pack_tiles_f32(static_cast<const float*>(W), N, taps, C, W_F16, img, pw, err);
    else if (w_ggml_type == -GGML_Q8_0) ok = pack_tiles_f32(static_cast<const float*>(W), N, taps, C, W_Q8_0, img, pw, err);
    else ok = taps == 1 && pack_tiles_ggml(W, w_ggml_type, N, C, img, pw, err);
    if (!ok) return set_error(ctx, ACE_GGML_ERR_INVALID_ARG, err.empty() ? "bad weight spec" : err);

    TmpAllocs tmp;
    const size_t a_elems = (size_t)B * L * C;
    float* dA32 = static_cast<float*>(tmp.get(a_elems * 4));
    __half* dA = static_cast<__half*>(tmp.get(a_elems * 2));
    pw.dev = tmp.get(img.size());
    float* dOut = static_cast<float*>(tmp.get(out_elems * 4));
    __half* dOut16 = static_cast<__half*>(tmp.get(out_elems * 2));
    float* dBias = bias ? static_cast<float*>(tmp.get((size_t)N * 4)) : nullptr;
    if (!dA32 || !dA || !pw.dev || !dOut || !dOut16 || (bias && !dBias)) return cuda_fail(ctx, "debug_gemm cudaMalloc failed", cudaGetLastError());
    cudaStream_t st = ctx->stream;
    cudaMemcpyAsync(dA32, A, a_elems * 4, cudaMemcpyHostToDevice, st);
    cudaMemcpyAsync(pw.dev, img.data(), img.size(), cudaMemcpyHostToDevice, st);
    if (bias) cudaMemcpyAsync(dBias, bias, (size_t)N * 4, cudaMemcpyHostToDevice, st);
    if ((epilogue & 4) && resid) cudaMemcpyAsync(dOut, resid, out_elems * 4, cudaMemcpyHostToDevice, st);
    else cudaMemsetAsync(dOut, 0, out_elems * 4, st);
    launch_f32_to_f16(dA32, dA, (long)a_elems, st);
    if (pw.wtype != W_F16 && predecode_enabled()) {  // decoded twin, as the loaders build it (gemm.h policy picks the route)
        pw.dev_f16 = tmp.get(gemm_predecode_bytes(pw));
        if (!pw.dev_f16) return cuda_fail(ctx, "debug_gemm cudaMalloc failed", cudaGetLastError());
        cudaError_t pe = gemm_predecode(pw, pw.dev_f16, st);
        if (pe != cudaSuccess) return cuda_fail(ctx, "debug_gemm predecode failed", pe);
    }
    GemmProblem g;
    g.A = dA;
    g.B = B;
    g.L = L;
    g.C = C;
    g.a_lstride = C;
    g.a_bstride = (long)L * C;
    g.l_out = l_out;
    g.tap_stride = tap_stride;
    g.tap_offset = tap_offset;
    g.W = &pw;
    g.split_k = (epilogue & 8) ? 4 : 1;
    g.epi.bias = (epilogue & 1) ? dBias : nullptr;
    g.epi.ld_out = n_out;
    g.epi.out_bstride = (long)l_out * n_out;
    if (swiglu) {
        g.epi.mode = EPI_SWIGLU;
        g.epi.out_f16 = dOut16;
    } else {
        g.epi.out_f32 = dOut;
        if ((epilogue & 4) || (epilogue & 8)) g.epi.resid = dOut;  // accumulate on top of `resid` (or zeros)
    }
    cudaEventRecord(ctx->ev0, st);
    cudaError_t e = gemm_launch(g, st, simt != 0);
    cudaEventRecord(ctx->ev1, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return cuda_fail(ctx, "debug_gemm failed", e);
    if (swiglu) {
        std::vector<uint16_t> h(out_elems);
        cudaMemcpy(h.data(), dOut16, out_elems * 2, cudaMemcpyDeviceToHost);
        for (size_t i = 0; i < out_elems; ++i) out[i] = f16_to_f32(h[i]);
    } else {
        cudaMemcpy(out, dOut, out_elems * 4, cudaMemcpyDeviceToHost);
    }
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->last_device_ms = ms;
    return ACE_GGML_OK;
}

ace_ggml_status ace_b200_debug_attention(ace_ggml_context* ctx, const float* Q, const float* K, const float* V, const int32_t* key_mask,
                                         int32_t B, int32_t q_len, int32_t k_len, int32_t n_heads, int32_t n_kv, int32_t window, float* out,
                                         size_t out_size) {
    if (!ctx || !Q || !K || !V || !out || B <= 0 || q_len <= 0 || k_len <= 0 || n_heads <= 0 || n_kv <= 0 || n_heads % n_kv) return ACE_GGML_ERR_INVALID_ARG;
    DeviceGuard guard(ctx->device);
    const size_t nq = (size_t)B * q_len * n_heads * 128, nk = (size_t)B * k_len * n_kv * 128;
    if (out_size < nq * 4) return set_error(ctx, ACE_GGML_ERR_INVALID_ARG, "output buffer too small");
    TmpAllocs tmp;
    float* d32 = static_cast<float*>(tmp.get(std::max(nq, nk) * 4));
    __half* dQ = static_cast<__half*>(tmp.get(nq * 2));
    __half* dK = static_cast<__half*>(tmp.get(nk * 2));
    __half* dV = static_cast<__half*>(tmp.get(nk * 2));
    __half* dO = static_cast<__half*>(tmp.get(nq * 2));
    int32_t* dM = key_mask ? static_cast<int32_t*>(tmp.get((size_t)B * k_len * 4)) : nullptr;
    if (!d32 || !dQ || !dK || !dV || !dO || (key_mask && !dM)) return cuda_fail(ctx, "debug_attention cudaMalloc failed", cudaGetLastError());
    cudaStream_t st = ctx->stream;
    cudaMemcpyAsync(d32, Q, nq * 4, cudaMemcpyHostToDevice, st);
    launch_f32_to_f16(d32, dQ, (long)nq, st);
    cudaMemcpyAsync(d32, K, nk * 4, cudaMemcpyHostToDevice, st);
    launch_f32_to_f16(d32, dK, (long)nk, st);
    cudaMemcpyAsync(d32, V, nk * 4, cudaMemcpyHostToDevice, st);
    launch_f32_to_f16(d32, dV, (long)nk, st);
    if (key_mask) cudaMemcpyAsync(dM, key_mask, (size_t)B * k_len * 4, cudaMemcpyHostToDevice, st);
    cudaEventRecord(ctx->ev0, st);
    launch_attention(dQ, (long)n_heads * 128, dK, (long)n_kv * 128, dV, (long)n_kv * 128, dO, (long)n_heads * 128, dM, B, q_len, k_len,
                     n_heads, n_kv, window, st);
    cudaEventRecord(ctx->ev1, st);
    cudaError_t e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return cuda_fail(ctx, "debug_attention failed", e);
    std::vector<uint16_t> h(nq);
    cudaMemcpy(h.data(), dO, nq * 2, cudaMemcpyDeviceToHost);
    for (size_t i = 0; i < nq; ++i) out[i] = f16_to_f32(h[i]);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->last_device_ms = ms;
    return ACE_GGML_OK;
}

// CPU-only: open a weight file with the library's own readers and report the number of tensors (or -1 + message).
int64_t ace_b200_debug_probe_tensor_file(const char* path, int32_t is_gguf, char* err_out, size_t err_size) {
    if (!path) return -1;
    std::string err;
    std::unique_ptr<TensorFile> tf = is_gguf ? TensorFile::open_gguf(path, err) : TensorFile::open_safetensors(path, err);
    if (err_out && err_size) {
        std::snprintf(err_out, err_size, "%s", err.c_str());
    }
    if (!tf) return -1;
    std::vector<float> tmp;
    for (const TensorInfo& t : tf->tensors()) {  // every tensor must be decodable without leaving the mapping
        std::string e;
        if (t.numel() <= (1 << 22) && !tf->read_f32(t.name, tmp, nullptr, e)) {
            if (err_out && err_size) std::snprintf(err_out, err_size, "%s", e.c_str());
            return -1;
        }
    }
    return (int64_t)tf->tensors().size();
}

// CPU-only: quantise fp32 rows [N, K] to raw ggml rows of w_ggml_type (8 Q8_0 / 12 Q4_K / 14 Q6_K) with the library's own quantisers
ace_ggml_status ace_b200_debug_quantize_rows(const float* src, int32_t w_ggml_type, int32_t N, int32_t K, void* dst, size_t dst_size) {
    if (!src || !dst || N <= 0 || K <= 0) return ACE_GGML_ERR_INVALID_ARG;
    const size_t rb = ggml_row_bytes(w_ggml_type, K);
    if (rb == 0 || dst_size < rb * (size_t)N) return ACE_GGML_ERR_INVALID_ARG;
    for (int r = 0; r < N; ++r) {
        uint8_t* d = static_cast<uint8_t*>(dst) + (size_t)r * rb;
        if (w_ggml_type == GGML_Q8_0) quantize_row_q8_0(src + (size_t)r * K, d, K);
        else if (w_ggml_type == GGML_Q4_K) quantize_row_q4_k(src + (size_t)r * K, d, K);
        else if (w_ggml_type == GGML_Q6_K) quantize_row_q6_k(src + (size_t)r * K, d, K);
        else return ACE_GGML_ERR_UNSUPPORTED;
    }
    return ACE_GGML_OK;
}

}  // extern "C"
