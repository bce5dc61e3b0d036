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
