// ace_b200_* extensions (include/acestep_b200.h): batched forward, device-resident sampler, batched VAE decode,
// measurement hooks.  Same error / ownership conventions as the reference entry points in abi.cu.
#include "abi_internal.h"

#include <cstdio>

using namespace ace_abi;

extern "C" {

ace_ggml_status ace_b200_dit_forward_batched(ace_ggml_context* ctx, const float* hidden_states, const float* context_latents,
                                             const float* encoder_hidden_states, const int32_t* attention_mask,
                                             const int32_t* encoder_attention_mask, int32_t batch, int32_t seq_len, int32_t enc_len,
                                             const float* timesteps, const float* timesteps_r, float* out, size_t out_size) {
    if (!ctx || !out || !timesteps || !timesteps_r || seq_len <= 0 || batch <= 0) return ACE_GGML_ERR_INVALID_ARG;
    if (!ctx->dit_loaded) return set_error(ctx, ACE_GGML_ERR, "dit not loaded");
    if (enc_len < 0) enc_len = 0;
    return dit_forward_impl(ctx, hidden_states, context_latents, encoder_hidden_states, attention_mask, encoder_attention_mask, batch,
                            seq_len, enc_len, timesteps, timesteps_r, out, out_size, "ace_b200_dit_forward_batched");
}

ace_ggml_status ace_b200_generate_latents(ace_ggml_context* ctx, const float* noise, const float* context_latents,
                                          const float* encoder_hidden_states, const float* uncond_encoder_hidden_states,
                                          const int32_t* attention_mask, const int32_t* encoder_attention_mask, int32_t batch,
                                          int32_t seq_len, int32_t enc_len, const float* schedule, int32_t n_steps, float guidance,
                                          float* out_latents, size_t out_size) {
    if (!ctx || !noise || !schedule || !out_latents || batch <= 0 || seq_len <= 0 || n_steps <= 0) return ACE_GGML_ERR_INVALID_ARG;
    if (!ctx->dit_loaded) return set_error(ctx, ACE_GGML_ERR, "dit not loaded");
    if (enc_len < 0) enc_len = 0;
    DeviceGuard guard(ctx->device);
    const size_t needed = (size_t)batch * seq_len * ctx->dit.cfg.audio_acoustic_hidden_dim * 4;
    if (out_size < needed) return set_error(ctx, ACE_GGML_ERR_INVALID_ARG, "output buffer too small");
    ace_ggml_status s = sample_latents_device(ctx, noise, context_latents, encoder_hidden_states, uncond_encoder_hidden_states,
                                              attention_mask, encoder_attention_mask, batch, seq_len, enc_len, schedule, n_steps, guidance);
    if (s != ACE_GGML_OK) return s;
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) return cuda_fail(ctx, "sampler failed", e);
    e = cudaMemcpy(out_latents, ctx->d_x, needed, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return cuda_fail(ctx, "latent download failed", e);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->last_device_ms = ms;
    return ACE_GGML_OK;
}

ace_ggml_status ace_b200_vae_decode_batched(ace_ggml_context* ctx, const float* latents, int32_t batch, int32_t n_frames, float* out,
                                            size_t out_size) {
    if (!ctx || !latents || !out || n_frames <= 0 || batch <= 0) return ACE_GGML_ERR_INVALID_ARG;
    if (!ctx->vae_loaded) return set_error(ctx, ACE_GGML_ERR, "vae not loaded");
    return vae_decode_impl(ctx, latents, batch, n_frames, out, out_size);
}

ace_ggml_status ace_b200_generate_audio(ace_ggml_context* ctx, const float* noise, const float* context_latents,
                                        const float* encoder_hidden_states, const float* uncond_encoder_hidden_states,
                                        const int32_t* attention_mask, const int32_t* encoder_attention_mask, int32_t batch,
                                        int32_t seq_len, int32_t enc_len, const float* schedule, int32_t n_steps, float guidance,
                                        float* out_latents, float* out_audio, size_t out_audio_size) {
    if (!ctx || !noise || !schedule || !out_audio || batch <= 0 || seq_len <= 0 || n_steps <= 0) return ACE_GGML_ERR_INVALID_ARG;
    if (!ctx->dit_loaded) return set_error(ctx, ACE_GGML_ERR, "dit not loaded");
    if (!ctx->vae_loaded) return set_error(ctx, ACE_GGML_ERR, "vae not loaded");
    if (ctx->vae.cfg.decoder_input_channels != ctx->dit.cfg.audio_acoustic_hidden_dim)
        return set_error(ctx, ACE_GGML_ERR, "DiT latent width does not match the VAE input channels");
    if (enc_len < 0) enc_len = 0;
    DeviceGuard guard(ctx->device);
    const VaeConfig& vc = ctx->vae.cfg;
    const size_t a_bytes = (size_t)batch * seq_len * vc.hop_length * vc.audio_channels * 4;
    if (out_audio_size < a_bytes) return set_error(ctx, ACE_GGML_ERR_INVALID_ARG, "output buffer too small");
    if (!ctx->io.reserve(0, a_bytes)) return set_error(ctx, ACE_GGML_ERR, "audio buffer allocation failed");
    ace_ggml_status s = sample_latents_device(ctx, noise, context_latents, encoder_hidden_states, uncond_encoder_hidden_states,
                                              attention_mask, encoder_attention_mask, batch, seq_len, enc_len, schedule, n_steps, guidance);
    if (s != ACE_GGML_OK) return s;
    const size_t l_bytes = (size_t)batch * seq_len * vc.decoder_input_channels * 4;
    std::string err;
    cudaError_t e = vae_decode_grouped(ctx, ctx->d_x, batch, seq_len, static_cast<float*>(ctx->io.d_out), err);
    cudaEventRecord(ctx->ev1, ctx->stream);
    if (e != cudaSuccess) return err.empty() ? cuda_fail(ctx, "vae decode launch failed", e) : set_error(ctx, ACE_GGML_ERR, err);
    e = cudaMemcpyAsync(out_audio, ctx->io.d_out, a_bytes, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess && out_latents) e = cudaMemcpyAsync(out_latents, ctx->d_x, l_bytes, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) return cuda_fail(ctx, "generate_audio failed", e);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->last_device_ms = ms;
    return ACE_GGML_OK;
}

ace_ggml_status ace_b200_create_on_device(const ace_ggml_init_params* params, int32_t device, ace_ggml_context** out_ctx) {
    if (device < 0) return ACE_GGML_ERR_INVALID_ARG;
    return create_context(params, device, out_ctx);
}

// ---- time-split VAE decode of one clip over several GPUs: the assembled clip lives on ONE device; every rank's final conv
// kernel stores its core samples straight into it (peer stores over NVLink / NVSwitch), so the split needs no gather pass.
ace_ggml_status ace_b200_peer_buffer_create(ace_ggml_context* ctx, size_t bytes, void* ipc_handle_out) {
    if (!ctx || bytes == 0) return ACE_GGML_ERR_INVALID_ARG;
    DeviceGuard guard(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (!(ctx->peer_kind == 1 && ctx->peer_bytes >= bytes)) {
        release_peer_buffer(ctx);
        void* p = nullptr;
        if (cudaMalloc(&p, bytes) != cudaSuccess) return cuda_fail(ctx, "peer buffer cudaMalloc failed", cudaGetLastError());
        ctx->peer_buf = static_cast<float*>(p);
        ctx->peer_bytes = bytes;
        ctx->peer_kind = 1;
    }
    if (ipc_handle_out) {
        cudaIpcMemHandle_t h;
        const cudaError_t e = cudaIpcGetMemHandle(&h, ctx->peer_buf);
        if (e != cudaSuccess) return cuda_fail(ctx, "cudaIpcGetMemHandle failed", e);
        static_assert(sizeof(h) == 64, "CUDA IPC handles are 64 bytes");
        std::memcpy(ipc_handle_out, &h, sizeof(h));
    }
    return ACE_GGML_OK;
}

ace_ggml_status ace_b200_peer_buffer_open(ace_ggml_context* ctx, const void* ipc_handle, size_t bytes) {
    if (!ctx || !ipc_handle || bytes == 0) return ACE_GGML_ERR_INVALID_ARG;
    DeviceGuard guard(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    release_peer_buffer(ctx);
    cudaIpcMemHandle_t h;
    std::memcpy(&h, ipc_handle, sizeof(h));
    void* p = nullptr;
    const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) return cuda_fail(ctx, "cudaIpcOpenMemHandle failed (no peer access between the two GPUs?)", e);
    ctx->peer_buf = static_cast<float*>(p);
    ctx->peer_bytes = bytes;
    ctx->peer_kind = 2;
    return ACE_GGML_OK;
}

ace_ggml_status ace_b200_peer_buffer_attach(ace_ggml_context* ctx, ace_ggml_context* owner) {
    if (!ctx || !owner || owner->peer_kind != 1 || !owner->peer_buf) return ACE_GGML_ERR_INVALID_ARG;
    DeviceGuard guard(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    release_peer_buffer(ctx);
    if (owner->device != ctx->device) {
        int can = 0;
        cudaDeviceCanAccessPeer(&can, ctx->device, owner->device);
        if (!can) return set_error(ctx, ACE_GGML_ERR_UNSUPPORTED, "no peer access between the two GPUs");
        const cudaError_t e = cudaDeviceEnablePeerAccess(owner->device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return cuda_fail(ctx, "cudaDeviceEnablePeerAccess failed", e);
        cudaGetLastError();
    }
    ctx->peer_buf = owner->peer_buf;
    ctx->peer_bytes = owner->peer_bytes;
    ctx->peer_kind = 3;
    return ACE_GGML_OK;
}

ace_ggml_status ace_b200_vae_decode_window(ace_ggml_context* ctx, const float* latents_window, int32_t n_window_frames,
                                           int32_t core_frame0, int32_t core_frames, int64_t dst_frame) {
    if (!ctx || !latents_window || n_window_frames <= 0 || core_frame0 < 0 || core_frames < 0 || core_frame0 + core_frames > n_window_frames ||
        dst_frame < 0)
        return ACE_GGML_ERR_INVALID_ARG;
    if (!ctx->vae_loaded) return set_error(ctx, ACE_GGML_ERR, "vae not loaded");
    if (!ctx->peer_buf) return set_error(ctx, ACE_GGML_ERR, "no peer buffer bound (ace_b200_peer_buffer_create / _open / _attach)");
    DeviceGuard guard(ctx->device);
    const VaeConfig& c = ctx->vae.cfg;
    const size_t end = (size_t)(dst_frame + core_frames) * c.hop_length * c.audio_channels * 4;
    if (end > ctx->peer_bytes) return set_error(ctx, ACE_GGML_ERR_INVALID_ARG, "output buffer too small");
    std::string err;
    if (!vae_workspace_reserve(ctx->vae, ctx->vws, 1, n_window_frames, err)) return set_error(ctx, ACE_GGML_ERR, err);
    cudaError_t e = cudaMemcpyAsync(ctx->vws.d_lat, latents_window, (size_t)n_window_frames * c.decoder_input_channels * 4, cudaMemcpyHostToDevice, ctx->stream);
    if (e != cudaSuccess) return cuda_fail(ctx, "latent upload failed", e);
    cudaEventRecord(ctx->ev0, ctx->stream);
    VaeOutWindow win{core_frame0, core_frame0 + core_frames, dst_frame};
    e = vae_decode_device(ctx->vae, ctx->vws, 1, n_window_frames, ctx->peer_buf, ctx->stream, &win);
    cudaEventRecord(ctx->ev1, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);  // the peer stores have landed when this returns
    if (e != cudaSuccess) return cuda_fail(ctx, "vae decode failed", e);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->last_device_ms = ms;
    return ACE_GGML_OK;
}

ace_ggml_status ace_b200_peer_buffer_read(ace_ggml_context* ctx, float* out, size_t bytes) {
    if (!ctx || !out) return ACE_GGML_ERR_INVALID_ARG;
    if (!ctx->peer_buf || bytes > ctx->peer_bytes) return set_error(ctx, ACE_GGML_ERR_INVALID_ARG, "output buffer too small");
    DeviceGuard guard(ctx->device);
    cudaError_t e = cudaMemcpyAsync(out, ctx->peer_buf, bytes, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) return cuda_fail(ctx, "audio download failed", e);
    return ACE_GGML_OK;
}

ace_ggml_status ace_b200_shift_schedule(float shift, int32_t n_steps, float* out) {
    if (!out || n_steps <= 0 || !(shift > 0.f)) return ACE_GGML_ERR_INVALID_ARG;
    for (int i = 0; i < n_steps; ++i) {
        const double u = 1.0 - (double)i / (double)n_steps;
        out[i] = (float)((double)shift * u / (1.0 + ((double)shift - 1.0) * u));
    }
    return ACE_GGML_OK;
}

double ace_b200_last_device_ms(const ace_ggml_context* ctx) { return ctx ? ctx->last_device_ms : 0.0; }
uint64_t ace_b200_kernel_launches(void) { return ace::g_kernel_launches.load(); }
const char* ace_b200_model_info(const ace_ggml_context* ctx_c) {
    if (!ctx_c) return "";
    ace_ggml_context* ctx = const_cast<ace_ggml_context*>(ctx_c);
    char buf[512];
    std::snprintf(buf, sizeof(buf), "device=%d dit=%s(%s, %.1f MB) vae=%s(%.1f MB)", ctx->device, ctx->dit_loaded ? "loaded" : "none",
                  ctx->dit_loaded ? ctx->dit.weight_desc.c_str() : "-", ctx->dit.weight_bytes / 1048576.0,
                  ctx->vae_loaded ? "loaded" : "none", ctx->vae.weight_bytes / 1048576.0);
    ctx->info = buf;
    return ctx->info.c_str();
}

void ace_b200_profile_enable(int32_t on) { ace::prof_enable(on != 0); }
void ace_b200_profile_collect(double* ms, double* flops, double* bytes, int64_t* launches) {
    ace::ProfTotals t;
    ace::prof_collect(t);
    for (int c = 0; c < ace::PROF_NCAT; ++c) {
        if (ms) ms[c] = t.ms[c];
        if (flops) flops[c] = t.flops[c];
        if (bytes) bytes[c] = t.bytes[c];
        if (launches) launches[c] = t.launches[c];
    }
}

}  // extern "C"
