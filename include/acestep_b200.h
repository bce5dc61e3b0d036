/* acestep_b200.h — additive extensions to the acestep_ggml C ABI (new symbols only; the reference symbols in
 * acestep_ggml.h keep their exact meaning).  They exist because the reference ABI evaluates one sample per call
 * from host buffers (acestep_ggml.h:96-108; the Python hook loops `for b in range(bsz)`,
 * scripts/run_non_ggml_real_case.py:516-527), which on a GPU would spend the step in PCIe copies:
 *
 *   ace_b200_dit_forward_batched   B samples per call (what the hook's loop becomes)
 *   ace_b200_generate_latents      device-resident Euler sampler: N steps, optional CFG, one H2D + one D2H
 *                                  (same update rule as acestep_ggml.cpp:2056-2086)
 *   ace_b200_vae_decode_batched    B clips per call, un-tiled (exact) decode
 *   ace_b200_generate_audio        sampler + VAE decode without leaving the device
 *   ace_b200_*_stats / debug_*     measurement + unit-test hooks (device time of the last call, launch count,
 *                                  direct access to the GEMM / attention kernels for parity tests)
 */
#ifndef ACESTEP_B200_H
#define ACESTEP_B200_H

#include "acestep_ggml.h"

#ifdef __cplusplus
extern "C" {
#endif

/* hidden_states [B, seq_len, 64], context_latents [B, seq_len, 128], encoder_hidden_states [B, enc_len, hidden],
 * attention_mask [B, seq_len] / encoder_attention_mask [B, enc_len] (nullable), timesteps / timesteps_r [B],
 * out [B, seq_len, 64]; out_size in bytes. */
ACE_GGML_API ace_ggml_status ace_b200_dit_forward_batched(
    ace_ggml_context * ctx, const float * hidden_states, const float * context_latents,
    const float * encoder_hidden_states, const int32_t * attention_mask, const int32_t * encoder_attention_mask,
    int32_t batch, int32_t seq_len, int32_t enc_len, const float * timesteps, const float * timesteps_r,
    float * out, size_t out_size);

/* Euler sampler on the device.  noise [B, seq_len, 64] is the initial x_T (the caller draws it, so any RNG —
 * including the reference's std::mt19937 sequence — can be reproduced); schedule [n_steps] are the t_i
 * (x -= v*(t_i - t_{i+1}); last step x -= v*t_i).  uncond_encoder_hidden_states == NULL disables CFG; otherwise
 * v = v_u + guidance*(v_c - v_u) with both branches evaluated in ONE batched forward of 2B samples.
 * out_latents [B, seq_len, 64]. */
ACE_GGML_API ace_ggml_status ace_b200_generate_latents(
    ace_ggml_context * ctx, const float * noise, const float * context_latents,
    const float * encoder_hidden_states, const float * uncond_encoder_hidden_states,
    const int32_t * attention_mask, const int32_t * encoder_attention_mask, int32_t batch, int32_t seq_len,
    int32_t enc_len, const float * schedule, int32_t n_steps, float guidance, float * out_latents, size_t out_size);

/* latents [B, n_frames, latent_ch] -> out [B, n_frames*hop, audio_ch] */
ACE_GGML_API ace_ggml_status ace_b200_vae_decode_batched(ace_ggml_context * ctx, const float * latents, int32_t batch,
                                                         int32_t n_frames, float * out, size_t out_size);

/* ace_b200_generate_latents followed by the VAE decode, latents never leave HBM. out_audio [B, seq_len*hop, audio_ch];
 * out_latents may be NULL. */
ACE_GGML_API ace_ggml_status ace_b200_generate_audio(
    ace_ggml_context * ctx, const float * noise, const float * context_latents,
    const float * encoder_hidden_states, const float * uncond_encoder_hidden_states,
    const int32_t * attention_mask, const int32_t * encoder_attention_mask, int32_t batch, int32_t seq_len,
    int32_t enc_len, const float * schedule, int32_t n_steps, float guidance, float * out_latents,
    float * out_audio, size_t out_audio_size);

/* The sampler the reference runs behind ace_ggml_generate_audio_* once the condition is encoded (static
 * ace_generate_audio_from_encoder, acestep_ggml.cpp:1901-2238), exposed for callers that already hold encoder states:
 * context_latents = [silence latent | chunk mask] (ACE_GGML_SILENCE_LATENT_F32 file, else VAE-encoded digital silence),
 * x_T from std::mt19937(seed) + std::normal_distribution<float> in frame-major order, the 8-step table nearest to `shift`,
 * device-resident Euler loop, VAE decode.  out_audio [seq_len*hop, audio_ch]. */
ACE_GGML_API ace_ggml_status ace_b200_generate_audio_from_encoder(
    ace_ggml_context * ctx, const float * encoder_hidden_states, const int32_t * encoder_attention_mask, int32_t enc_len,
    int32_t seq_len, float shift, int32_t seed, float * out_audio, size_t out_size, int32_t * out_audio_samples,
    int32_t * out_audio_channels);

/* The two condition encoders of the DiT checkpoint on their own (acestep_dit_model.cpp:1562-1739):
 * lyric: token embeddings [n_tokens, text_hidden_dim] -> [n_tokens, hidden]; timbre: reference-audio latents
 * [refer_len, timbre_hidden_dim] -> [hidden] (first time step of the encoded sequence). */
ACE_GGML_API ace_ggml_status ace_b200_lyric_encoder_forward(ace_ggml_context * ctx, const float * lyric_embeds,
                                                            int32_t n_tokens, float * out, size_t out_size);
ACE_GGML_API ace_ggml_status ace_b200_timbre_encoder_forward(ace_ggml_context * ctx, const float * refer_states,
                                                             int32_t refer_len, float * out, size_t out_size);

/* ace_ggml_create on an explicitly named CUDA device (ace_ggml_create itself takes ACE_B200_DEVICE, else the calling thread's
 * current device).  Contexts on different devices are independent and may be driven from different host threads. */
ACE_GGML_API ace_ggml_status ace_b200_create_on_device(const ace_ggml_init_params * params, int32_t device,
                                                       ace_ggml_context ** out_ctx);

/* Time-split VAE decode of ONE long clip over several GPUs (BASELINE configs[3]; replaces the reference's chunk / overlap loop,
 * acestep_ggml.cpp:2114-2229).  The assembled clip lives on one device:
 *   owner:  ace_b200_peer_buffer_create(ctx, bytes, handle[64])  allocates it and returns its CUDA IPC handle (handle may be NULL);
 *   other processes: ace_b200_peer_buffer_open(ctx, handle, bytes);   other contexts of the same process: ace_b200_peer_buffer_attach;
 *   every rank: ace_b200_vae_decode_window decodes its halo'd window [n_window_frames, latent_ch] and the LAST kernel of the decode
 *     stores the samples of window frames [core_frame0, core_frame0 + core_frames) at frame offset dst_frame of the clip — plain
 *     stores into peer memory over NVLink, no gather pass; the call returns when they have landed;
 *   owner, after the ranks have synchronised: ace_b200_peer_buffer_read copies the clip to the host. */
ACE_GGML_API ace_ggml_status ace_b200_peer_buffer_create(ace_ggml_context * ctx, size_t bytes, void * ipc_handle_out);
ACE_GGML_API ace_ggml_status ace_b200_peer_buffer_open(ace_ggml_context * ctx, const void * ipc_handle, size_t bytes);
ACE_GGML_API ace_ggml_status ace_b200_peer_buffer_attach(ace_ggml_context * ctx, ace_ggml_context * owner);
ACE_GGML_API ace_ggml_status ace_b200_vae_decode_window(ace_ggml_context * ctx, const float * latents_window,
                                                        int32_t n_window_frames, int32_t core_frame0, int32_t core_frames,
                                                        int64_t dst_frame);
ACE_GGML_API ace_ggml_status ace_b200_peer_buffer_read(ace_ggml_context * ctx, float * out, size_t bytes);

/* the reference's shift schedules (acestep_ggml.cpp:1484-1500): t_i = s*u/(1+(s-1)*u), u = 1 - i/n_steps.
 * For n_steps == 8 and shift in {1,2,3} this reproduces the three hard-coded tables. */
ACE_GGML_API ace_ggml_status ace_b200_shift_schedule(float shift, int32_t n_steps, float * out_schedule);

/* measurement: device time (CUDA events on the context's stream) of the kernels of the last successful call on this
 * context, excluding host<->device copies; and the number of kernels this library launched since process start. */
ACE_GGML_API double   ace_b200_last_device_ms(const ace_ggml_context * ctx);
ACE_GGML_API uint64_t ace_b200_kernel_launches(void);
ACE_GGML_API const char * ace_b200_model_info(const ace_ggml_context * ctx);
/* per-kernel-family device timing: enable(1) before a call, collect after it (then enable(0)).
 * families: 0 tap-GEMM, 1 attention, 2 element-wise / norm, 3 VAE output conv.  out arrays have 4 entries each. */
ACE_GGML_API void ace_b200_profile_enable(int32_t on);
ACE_GGML_API void ace_b200_profile_collect(double * ms, double * flops, double * bytes, int64_t * launches);

/* ---- allocation-free variant ------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 * ace_b200_attention_noalloc: the same attention kernel, arguments and results as ace_b200_debug_attention below
 * (Q [B, q_len, n_heads*128], K/V [B, k_len, n_kv*128] fp32, rounded to fp16 inside; key_mask [B, k_len] or NULL;
 * window < 0 = full; out fp32 [B, q_len, n_heads*128]; the pointers may be host or device memory), but its fp16 operand /
 * staging scratch is a static device array of ACE_B200_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds
 * until exit, and it calls no cudaMalloc / cudaFree.  Needs
 *   4*max(nq, nk) + 4*nq + 4*nk (+ 4*B*k_len with a mask) bytes,  nq = B*q_len*n_heads*128, nk = B*k_len*n_kv*128
 * (each piece rounded up to 256 B); a larger problem is rejected with ACE_GGML_ERR_INVALID_ARG and a message naming
 * the need and the limit.  All calls on one GPU share the array (it keeps the last call's contents; every call
 * overwrites what it reads and clears its output region first): calls that overlap on the same GPU race. */
#define ACE_B200_NOALLOC_SCRATCH_BYTES ((size_t)16 << 20)
ACE_GGML_API ace_ggml_status ace_b200_attention_noalloc(ace_ggml_context * ctx, const float * Q, const float * K,
                                                        const float * V, const int32_t * key_mask, int32_t B,
                                                        int32_t q_len, int32_t k_len, int32_t n_heads, int32_t n_kv,
                                                        int32_t window, float * out, size_t out_size);

/* ---- unit-test hooks (tests/ only) -------------------------------------------------------------------------
 * debug_gemm: D[B, l_out, N] = tapGEMM(A fp32 [B, L, C] rounded to fp16, W) through the tensor-core kernel (simt = 0)
 * or the plain-FMA debug kernel (simt = 1).  W is given either as fp32 [N, taps, C] (w_ggml_type = 0 -> F16 tiles,
 * 8 -> quantised to Q8_0 on the fly) or as raw ggml rows of w_ggml_type (8 Q8_0 / 12 Q4_K / 14 Q6_K), taps = 1.
 * epilogue: bit 0 bias[N], bit 1 SwiGLU over interleaved rows (out [.., N/2]), bit 2 residual add of `resid`,
 * bit 3 split-K = 4.  Output is fp32 (SwiGLU: the fp16 result widened). */
ACE_GGML_API ace_ggml_status ace_b200_debug_gemm(ace_ggml_context * ctx, const float * A, int32_t B, int32_t L,
                                                 int32_t C, const void * W, int32_t w_ggml_type, int32_t N,
                                                 int32_t taps, int32_t tap_stride, int32_t tap_offset, int32_t l_out,
                                                 const float * bias, const float * resid, int32_t epilogue,
                                                 int32_t simt, float * out, size_t out_size);
/* debug_attention: Q [B, q_len, n_heads*128], K/V [B, k_len, n_kv*128] fp32 (rounded to fp16 inside), window < 0 = full */
ACE_GGML_API ace_ggml_status ace_b200_debug_attention(ace_ggml_context * ctx, const float * Q, const float * K,
                                                      const float * V, const int32_t * key_mask, int32_t B,
                                                      int32_t q_len, int32_t k_len, int32_t n_heads, int32_t n_kv,
                                                      int32_t window, float * out, size_t out_size);
/* CPU-only: pack raw ggml rows (or fp32 when w_ggml_type == 0) into device tiles and decode them back with the host
 * emulation of the kernel's decoder; `ref` receives the ggml dequantisation of the same rows. No GPU needed. */
ACE_GGML_API ace_ggml_status ace_b200_debug_pack_roundtrip(const void * W, int32_t w_ggml_type, int32_t N, int32_t K,
                                                           float * unpacked, float * ref);

/* CPU-only: fp32 rows [N, K] -> raw ggml rows of w_ggml_type (8 Q8_0 / 12 Q4_K / 14 Q6_K) with the library's own quantisers, the ones
 * ACE_GGML_*_WEIGHT_QTYPE runs at load; bit-identical to ggml_quantize_chunk (tests/test_wpack_cpu.py). */
ACE_GGML_API ace_ggml_status ace_b200_debug_quantize_rows(const float * src, int32_t w_ggml_type, int32_t N, int32_t K,
                                                          void * dst, size_t dst_size);

/* CPU-only: parse a .gguf (is_gguf != 0) or .safetensors file with the library's readers and decode every small tensor; returns the
 * tensor count, or -1 with the reader's message in err_out.  Used by the malformed-file tests. */
ACE_GGML_API int64_t ace_b200_debug_probe_tensor_file(const char * path, int32_t is_gguf, char * err_out, size_t err_size);

#ifdef __cplusplus
}
#endif
#endif /* ACESTEP_B200_H */
