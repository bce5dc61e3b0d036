/* acestep_ggml.h — C ABI of the B200-native DiT-denoise + VAE-decode library.
 *
 * Drop-in boundary: every symbol below has the name, argument order, argument
 * meaning and status convention of the reference library's public header
 * (reference: acestep_ggml/cpp/acestep_ggml.h:21-152), so a caller that
 * dlopen()s / links `libacestep_ggml` can be pointed at `libacestep_b200.so`
 * without source changes.  Only plain pointers and sizes cross the boundary.
 *
 *   hot path:   ace_ggml_create / destroy / last_error, ace_ggml_load_dit,
 *     ace_ggml_load_vae, ace_ggml_vae_get_info, ace_ggml_dit_forward, ace_ggml_vae_decode
 *   callers either side (§8f):    ace_ggml_load_text_encoder / load_lm, the four
 *     ace_ggml_text_encoder_forward* entry points, ace_ggml_vae_encode and the three
 *     ace_ggml_generate_audio_* wrappers (csrc/abi_gen.cu)
 *   Every entry point runs on sm_100a; none has a CPU fallback (a missing device or a
 *   failed launch is ACE_GGML_ERR with the reason in ace_ggml_last_error).
 *
 * Additive extensions (batched / device-resident sampler) live in acestep_b200.h.
 */
#ifndef ACESTEP_GGML_H
#define ACESTEP_GGML_H

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#  define ACE_GGML_API __attribute__((visibility("default")))
#else
#  define ACE_GGML_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* opaque handle (reference: acestep_ggml.h:21) */
typedef struct ace_ggml_context ace_ggml_context;

/* reference: acestep_ggml.h:23-29 */
typedef enum ace_ggml_status {
    ACE_GGML_OK              = 0,
    ACE_GGML_ERR             = 1,
    ACE_GGML_ERR_INVALID_ARG = 2,
    ACE_GGML_ERR_IO          = 3,
    ACE_GGML_ERR_UNSUPPORTED = 4
} ace_ggml_status;

/* reference: acestep_ggml.h:31-35.  n_threads and use_metal are accepted and
 * ignored (there is no CPU or Metal path); compute_buffer_bytes is a hint. */
typedef struct ace_ggml_init_params {
    int32_t n_threads;
    int32_t use_metal;
    size_t  compute_buffer_bytes;
} ace_ggml_init_params;

/* lifecycle — reference: acestep_ggml.h:37-39, acestep_ggml.cpp:192-236 */
ACE_GGML_API ace_ggml_status ace_ggml_create(const ace_ggml_init_params * params, ace_ggml_context ** out_ctx);
ACE_GGML_API void            ace_ggml_destroy(ace_ggml_context * ctx);
ACE_GGML_API const char *    ace_ggml_last_error(const ace_ggml_context * ctx);

/* model load — reference: acestep_ggml.h:41-44, acestep_ggml.cpp:260-357 (DiT), 529-632 (VAE) */
ACE_GGML_API ace_ggml_status ace_ggml_load_lm(ace_ggml_context * ctx, const char * model_dir);
ACE_GGML_API ace_ggml_status ace_ggml_load_text_encoder(ace_ggml_context * ctx, const char * model_dir);
ACE_GGML_API ace_ggml_status ace_ggml_load_dit(ace_ggml_context * ctx, const char * model_dir);
ACE_GGML_API ace_ggml_status ace_ggml_load_vae(ace_ggml_context * ctx, const char * model_dir);

/* VAE — reference: acestep_ggml.h:45-61, acestep_ggml.cpp:634-974.
 * latents [n_frames, latent_channels] f32 time-major; out [n_frames*hop, audio_channels] f32
 * interleaved; out_size in BYTES. */
ACE_GGML_API ace_ggml_status ace_ggml_vae_get_info(ace_ggml_context * ctx, int32_t * latent_channels,
                                                   int32_t * audio_channels, int32_t * hop_length);
ACE_GGML_API ace_ggml_status ace_ggml_vae_decode(ace_ggml_context * ctx, const float * latents, int32_t n_frames,
                                                 float * out, size_t out_size);
ACE_GGML_API ace_ggml_status ace_ggml_vae_encode(ace_ggml_context * ctx, const float * audio, int32_t n_samples,
                                                 float * out, size_t out_size);

/* text encoder (Qwen3 stack, csrc/abi_gen.cu) — reference: acestep_ggml.h:63-94 */
ACE_GGML_API ace_ggml_status ace_ggml_text_encoder_forward(ace_ggml_context * ctx, const int32_t * token_ids,
                                                           int32_t n_tokens, float * out, size_t out_size);
ACE_GGML_API ace_ggml_status ace_ggml_text_encoder_forward_masked(ace_ggml_context * ctx, const int32_t * token_ids,
                                                                  const int32_t * attention_mask, int32_t n_tokens,
                                                                  float * out, size_t out_size);
ACE_GGML_API ace_ggml_status ace_ggml_text_encoder_forward_embeddings(ace_ggml_context * ctx,
                                                                      const int32_t * token_ids, int32_t n_tokens,
                                                                      float * out, size_t out_size);
ACE_GGML_API ace_ggml_status ace_ggml_text_encoder_forward_layers(ace_ggml_context * ctx, const int32_t * token_ids,
                                                                  const int32_t * attention_mask, int32_t n_tokens,
                                                                  int32_t n_layers, int32_t apply_final_norm,
                                                                  float * out, size_t out_size);

/* one evaluation of the DiT velocity network — reference: acestep_ggml.h:96-108,
 * acestep_ggml.cpp:1304-1482, acestep_dit_model.cpp:1316-1560.
 *   hidden_states [seq_len, 64], context_latents [seq_len, 128], encoder_hidden_states [enc_len, hidden],
 *   attention_mask [seq_len] / encoder_attention_mask [enc_len] int32 (non-zero = keep; NULL = all ones),
 *   out [seq_len, 64] f32, out_size in BYTES. */
ACE_GGML_API ace_ggml_status ace_ggml_dit_forward(ace_ggml_context * ctx, const float * hidden_states,
                                                  const float * context_latents,
                                                  const float * encoder_hidden_states,
                                                  const int32_t * attention_mask,
                                                  const int32_t * encoder_attention_mask, int32_t seq_len,
                                                  int32_t enc_len, float timestep, float timestep_r, float * out,
                                                  size_t out_size);

/* token-to-audio wrappers (text / lyric / timbre encoders -> seed-exact noise -> sampler -> VAE) —
 * reference: acestep_ggml.h:110-152 */
ACE_GGML_API ace_ggml_status ace_ggml_generate_audio_simple(ace_ggml_context * ctx, const int32_t * token_ids,
                                                            int32_t n_tokens, int32_t seq_len, float shift,
                                                            int32_t seed, float * out_audio, size_t out_size,
                                                            int32_t * out_audio_samples,
                                                            int32_t * out_audio_channels);
ACE_GGML_API ace_ggml_status ace_ggml_generate_audio_style_lyric_simple(
    ace_ggml_context * ctx, const int32_t * style_token_ids, int32_t n_style_tokens,
    const int32_t * lyric_token_ids, int32_t n_lyric_tokens, int32_t seq_len, float shift, int32_t seed,
    float * out_audio, size_t out_size, int32_t * out_audio_samples, int32_t * out_audio_channels);
ACE_GGML_API ace_ggml_status ace_ggml_generate_audio_style_lyric_timbre_simple(
    ace_ggml_context * ctx, const int32_t * style_token_ids, int32_t n_style_tokens,
    const int32_t * lyric_token_ids, int32_t n_lyric_tokens,
    const float * refer_audio_acoustic_hidden_states, const int32_t * refer_audio_order_mask,
    int32_t n_refer_audio, int32_t refer_audio_len, int32_t seq_len, float shift, int32_t seed,
    float * out_audio, size_t out_size, int32_t * out_audio_samples, int32_t * out_audio_channels);

#ifdef __cplusplus
}
#endif
#endif /* ACESTEP_GGML_H */
