// Launchers of the non-GEMM kernels of the path (element-wise / normalisation / attention / VAE tail).
// Every function enqueues on `st` and returns; errors surface through cudaGetLastError() at the call site.
#pragma once
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <atomic>

namespace ace {

extern std::atomic<uint64_t> g_kernel_launches;
inline void note_launch() { g_kernel_launches.fetch_add(1, std::memory_order_relaxed); }

// Optional per-kernel-family device timing (off by default; bench.py turns it on for ONE extra pass so the timed
// pass carries no event overhead): every launcher brackets its kernel with a CUDA event pair on the launching stream.
enum ProfCat : int { PROF_GEMM = 0, PROF_ATTN = 1, PROF_ELEM = 2, PROF_VAE_TAIL = 3, PROF_NCAT = 4 };
struct ProfTotals {
    double ms[PROF_NCAT];
    double flops[PROF_NCAT];   // algorithmic FLOPs
    double bytes[PROF_NCAT];   // algorithmic HBM bytes (where the launcher knows them)
    long launches[PROF_NCAT];
};
bool prof_enabled();
void prof_enable(bool on);                     // clears previous records
void prof_begin(int cat, double flops, double bytes, cudaStream_t st);
void prof_end(cudaStream_t st);
void prof_collect(ProfTotals& out);            // synchronises on the recorded events
struct ProfScope {
    cudaStream_t st;
    bool on;
    ProfScope(int cat, double flops, double bytes, cudaStream_t s) : st(s), on(prof_enabled()) {
        if (on) prof_begin(cat, flops, bytes, st);
    }
    ~ProfScope() {
        if (on) prof_end(st);
    }
};

// Function attributes (MaxDynamicSharedMemorySize), SM counts and occupancy answers are per DEVICE, not per process: a second
// context on another device of the same process must set / query them again.  `DeviceOnce` runs `fn` once per device
// (thread-safe; the fast path is one relaxed atomic load) and caches one int per device.
constexpr int ACE_MAX_DEVICES = 64;
struct DeviceOnce {
    std::atomic<int> done[ACE_MAX_DEVICES];
    int value[ACE_MAX_DEVICES];
    std::atomic_flag lock = ATOMIC_FLAG_INIT;
    DeviceOnce() {
        for (auto& d : done) d.store(0, std::memory_order_relaxed);
        for (auto& v : value) v = 0;
    }
    // fn(int& value) -> cudaError_t ; the result is remembered only on success
    template <typename F>
    cudaError_t run(F&& fn, int* out_value = nullptr) {
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev < 0 || dev >= ACE_MAX_DEVICES) return cudaErrorInvalidDevice;
        cudaError_t e = cudaSuccess;
        if (!done[dev].load(std::memory_order_acquire)) {
            while (lock.test_and_set(std::memory_order_acquire)) {}
            if (!done[dev].load(std::memory_order_relaxed)) {
                e = fn(value[dev]);
                if (e == cudaSuccess) done[dev].store(1, std::memory_order_release);
            }
            lock.clear(std::memory_order_release);
        }
        if (out_value) *out_value = value[dev];
        return e;
    }
};

// Launch with the programmatic-stream-serialization attribute (PDL): the kernel may start while its predecessor in the
// stream drains; every kernel launched this way executes ptx::pdl_wait() before its first global access.
// ACE_B200_PDL=0 turns the attribute off (plain stream order).
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ---- DiT input side ------------------------------------------------------------------------------
// x0[b, p, k*in_ch + c] = (c < ctx_dim ? context[b, 2p+k, c] : hidden[b, 2p+k, c-ctx_dim]), zero past T
// reference: acestep_dit_model.cpp:1343-1377 (pack + zero pad to a patch multiple)
void launch_pack_dit_input(const float* hidden, const float* context, __half* x0, int B, int T, int S, int patch,
                           int ctx_dim, int audio_dim, cudaStream_t st);
void launch_f32_to_f16(const float* src, __half* dst, long n, cudaStream_t st);
// patch_mask[b, p] = any(mask[b, p*patch + k] != 0); mask == nullptr -> all ones  (acestep_dit_model.cpp:1433-1449)
void launch_patch_mask(const int32_t* mask, int32_t* patch_mask, int B, int T, int S, int patch, cudaStream_t st);
// 256-d sinusoidal features of 1000*t: [cos(0..127) | sin(0..127)], freqs exp(-ln(1e4) i/128)
// reference: acestep_dit_model.cpp:1261-1284.  t: [n] device floats -> out fp16 [n, 256]
void launch_timestep_features(const float* t, __half* out, int n, cudaStream_t st);
// mod[b, layer, j, :] = table[layer][j, :] + proj[b, j, :]   with +1 folded into the scale rows (j = 1, 4)
// out_mod[b, j, :]    = out_table[j, :] + temb_a[b, :] + temb_b[b, :]        (+1 on j = 1)
// reference: acestep_dit_model.cpp:1469-1481, 1522-1526, 1537-1549
void launch_build_modulation(const float* tables /*[layers,6,H]*/, const float* proj /*[B,6,H]*/, float* mod,
                             const float* out_table /*[2,H]*/, const float* temb_a, const float* temb_b,
                             float* out_mod /*[B,2,H]*/, int B, int layers, int H, cudaStream_t st, bool shared_t = false);
// shared_t: every sample of the batch is at the same timestep, proj / temb_a / temb_b hold ONE row that all B samples read

// ---- normalisation -------------------------------------------------------------------------------
// y = fp16( x / sqrt(mean(x^2)+eps) * w * scale1[b] + shift[b] );  scale1/shift may be null (plain RMSNorm)
// reference: ggml_rms_norm (ggml-cpu/ops.cpp:3646-3697) + acestep_dit_model.cpp:1097-1106, 1477-1481
// y32 (optional) receives the same result unrounded (final norms of the condition / text encoders, whose output is fp32)
void launch_rmsnorm_mod(const float* x, const float* w, const float* scale1, const float* shift, long mod_bstride,
                        __half* y, int B, int rows_per_b, int H, float eps, cudaStream_t st, float* y32 = nullptr,
                        bool early_mod = false);  // early_mod: scale1 / shift may be read before the PDL wait (see the kernel)
// out[i, :] = table[ids[i], :]  (fp32 rows; ggml_get_rows + cast, qwen_model.cpp:553-554); ids outside [0, vocab) give zeros
void launch_gather_rows(const float* table, const int32_t* ids, float* out, int n, int width, int vocab, cudaStream_t st);
// dst[l, 0..c_dst) = fp16(src[l, 0..c_src)) zero-padded (VAE encoder input: 2 audio channels -> 8, acestep_vae_model.cpp:1018-1027)
void launch_pad_channels_f16(const float* src, __half* dst, long rows, int c_src, int c_dst, cudaStream_t st);
// dst[l, 0..c_dst) = src[l, 0..c_dst) of rows c_src wide (posterior mean = first half of the encoder's channels, :1039-1043)
void launch_take_channels_f32(const float* src, float* dst, long rows, int c_src, int c_dst, cudaStream_t st);
// in place on a token-major fp16 buffer: per (token, head) RMSNorm over head_dim=128 with weight w, then NEOX RoPE
// (pairs (i, i+64), angle pos * theta^(-2i/128)) when rope_tab != null.  rope_tab: [S, 64] (cos, sin) float2.
// reference: acestep_dit_model.cpp:1202-1210, ggml_rope_ext NEOX (ggml-cpu/ops.cpp:5637-5777)
void launch_headnorm_rope(__half* buf, long row_stride, int col0, int n_heads, const float* w, const float2* rope_tab,
                          int B, int rows_per_b, float eps, cudaStream_t st);
void launch_build_rope_table(float2* tab, int S, int half_dim, float theta, cudaStream_t st);

// ---- attention -----------------------------------------------------------------------------------
// O[b, q, h, :] = softmax_k( Q[b,q,h,:].K[b,k,h/rep,:] / sqrt(128) + mask ) V      (head_dim = 128)
// Q/K/V/O are token-major fp16 with independent row strides; key k is visible to query q iff
// (window < 0 or |q-k| <= window) and (key_mask == null or key_mask[b,k] != 0) and (not causal or k <= q).
// reference: acestep_dit_model.cpp:1132-1173 (mask), 1212-1256 (GQA repeat, scale, softmax, PV); qwen_model.cpp:603-624 (causal)
void launch_attention(const __half* Q, long q_stride, const __half* K, long k_stride, const __half* V, long v_stride,
                      __half* O, long o_stride, const int32_t* key_mask, int B, int q_len, int k_len, int n_heads,
                      int n_kv_heads, int window, cudaStream_t st, bool causal = false);

// tcgen05 / TMEM / TMA implementation (attention_tc.cu), the default; ACE_B200_ATTN=mma selects the mma.sync kernel of
// attention.cu (kept as the independent implementation the tensor-memory kernel is unit-tested against)
bool attention_use_tc();
cudaError_t launch_attention_tc(const __half* Q, long q_stride, const __half* K, long k_stride, const __half* V, long v_stride,
                                __half* O, long o_stride, const int32_t* key_mask, int B, int q_len, int k_len, int n_heads,
                                int n_kv_heads, int window, cudaStream_t st);

// ---- sampler -------------------------------------------------------------------------------------
// x -= v * dt   (acestep_ggml.cpp:2075-2085);  with CFG: v = v_u + g (v_c - v_u)
void launch_euler_step(float* x, const float* v_cond, const float* v_uncond, float guidance, float dt, long n,
                       cudaStream_t st);

// ---- VAE tail ------------------------------------------------------------------------------------
// final conv: out[b, l, co] = sum_{tap, c} xs[b, l + tap - pad, c] * w[co, tap, c]  (+bias) for tiny C_out (audio channels)
// reference: acestep_vae_model.cpp:999-1000 (conv2 k7, no bias)
// Only positions l in [l_lo, l_hi) are computed and stored, at out[b*out_bstride + l + out_shift] (defaults: the whole clip in place).
// A time-split decode passes its core range and a pointer into the assembled clip — which may be PEER memory of another GPU: the
// stores then travel over NVLink and the gather of the split costs no extra pass.
void launch_conv_small_out(const __half* xs, const __half* w /*[Cout, taps, C]*/, const float* bias, float* out, int B,
                           long L, int C, int Cout, int taps, int pad, cudaStream_t st, long l_lo = 0, long l_hi = -1,
                           long out_shift = 0, long out_bstride = -1);

}  // namespace ace
