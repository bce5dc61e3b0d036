// The "next" rows of the path behind the reference's own entry points: VAE encode, the Qwen3 text encoder, the
// lyric / timbre condition encoders + text projector + sequence packing, and the self-contained sampler
// (ace_ggml_generate_audio_*): context latents, seed-exact noise, shift schedule, device-resident Euler loop, VAE decode.
// reference: acestep_ggml.cpp:975-1072 (vae_encode), 1074-1302 (text encoder entry points), 1484-1573 (schedule, silence
// latent file), 1625-1899 (projection, lyric / timbre encoders, packing), 1901-2238 (ace_generate_audio_from_encoder),
// 2240-2644 (generate_audio_*)
#include "abi_internal.h"

#include <cmath>
#include <cstdio>
#include <fstream>
#include <numeric>
#include <random>

using namespace ace_abi;

namespace {

int read_nonneg_env_int(const char* key, int fallback) {  // acestep_ggml.cpp:1502-1511
    if (const char* v = std::getenv(key)) {
        char* endp = nullptr;
        const long long x = std::strtoll(v, &endp, 10);
        if (endp && endp != v && x >= 0) return (int)x;
    }
    return fallback;
}

// one of the three hard-coded 8-step tables, picked by the nearest shift in {1, 2, 3} (acestep_ggml.cpp:1484-1500)
void pick_shift_schedule(float shift, std::vector<float>& out) {
    const float d1 = std::fabs(shift - 1.0f), d2 = std::fabs(shift - 2.0f), d3 = std::fabs(shift - 3.0f);
    const float s = (d1 <= d2 && d1 <= d3) ? 1.0f : ((d2 <= d1 && d2 <= d3) ? 2.0f : 3.0f);
    out.resize(8);
    ace_b200_shift_schedule(s, 8, out.data());
}

// flat fp32 file of [frames, dim]; frame t of the result = frame min(t, frames-1) of the file (acestep_ggml.cpp:1513-1573)
bool load_silence_latent_f32(const char* path, int seq_len, int dim, std::vector<float>& out, std::string& err) {
    if (!path || !path[0]) { err = "empty silence latent path"; return false; }
    std::ifstream fin(path, std::ios::binary | std::ios::ate);
    if (!fin) { err = std::string("failed to open silence latent file: ") + path; return false; }
    const std::streamsize size = fin.tellg();
    if (size <= 0) { err = "silence latent file is empty"; return false; }
    if (size % 4 != 0) { err = "silence latent file has invalid byte size"; return false; }
    const size_t n_floats = (size_t)size / 4;
    if (n_floats % (size_t)dim != 0) { err = "silence latent float count is not divisible by latent dim"; return false; }
    const size_t n_frames = n_floats / (size_t)dim;
    if (n_frames == 0) { err = "silence latent has zero frames"; return false; }
    std::vector<float> full(n_floats);
    fin.seekg(0, std::ios::beg);
    fin.read(reinterpret_cast<char*>(full.data()), size);
    if (!fin) { err = "failed to read silence latent file"; return false; }
    out.assign((size_t)seq_len * dim, 0.f);
    for (int t = 0; t < seq_len; ++t)
        std::memcpy(out.data() + (size_t)t * dim, full.data() + std::min<size_t>((size_t)t, n_frames - 1) * dim, (size_t)dim * 4);
    return true;
}

// ---------------------------------------------------------------- device helpers
ace_ggml_status vae_encode_impl(ace_ggml_context* ctx, const float* audio, int n_samples, float* out, size_t out_size) {
    DeviceGuard guard(ctx->device);
    const VaeConfig& c = ctx->vae.cfg;
    // the reference checks the floor(n_samples / hop) size first, then the actual one (acestep_ggml.cpp:990-995, 1053-1058)
    const size_t expected = (size_t)(n_samples / c.hop_length) * c.decoder_input_channels * sizeof(float);
    if (out_size < expected) return set_error(ctx, ACE_GGML_ERR_INVALID_ARG, "output buffer too small");
    if (!ctx->vae.has_encoder) return set_error(ctx, ACE_GGML_ERR, "vae encode failed");
    const long frames = vae_encode_out_frames(ctx->vae, n_samples);
    if (frames <= 0) return set_error(ctx, ACE_GGML_ERR, "unexpected vae output shape");
    const size_t needed = (size_t)frames * c.decoder_input_channels * sizeof(float);
    if (out_size < needed) return set_error(ctx, ACE_GGML_ERR_INVALID_ARG, "output buffer too small for actual output");
    std::string err;
    if (!vae_encode_workspace_reserve(ctx->vae, ctx->vews, n_samples, err)) return set_error(ctx, ACE_GGML_ERR, err);
    const size_t in_bytes = (size_t)n_samples * c.audio_channels * 4;
    if (!ctx->io.reserve(in_bytes, needed)) return set_error(ctx, ACE_GGML_ERR, "VAE I/O buffer allocation failed");
    cudaError_t e = cudaMemcpyAsync(ctx->io.d_in, audio, in_bytes, cudaMemcpyHostToDevice, ctx->stream);
    if (e != cudaSuccess) return cuda_fail(ctx, "audio upload failed", e);
    cudaEventRecord(ctx->ev0, ctx->stream);
    e = vae_encode_device(ctx->vae, ctx->vews, static_cast<const float*>(ctx->io.d_in), n_samples, static_cast<float*>(ctx->io.d_out), ctx->stream);
    cudaEventRecord(ctx->ev1, ctx->stream);
    if (e != cudaSuccess) return cuda_fail(ctx, "vae encode launch failed", e);
    e = cudaMemcpyAsync(out, ctx->io.d_out, needed, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) return cuda_fail(ctx, "graph compute failed", e);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->last_device_ms = ms;
    return ACE_GGML_OK;
}

// small device scratch owned by the context for the condition side: ids, masks, fp32 staging rows
struct CondScratch {
    ace_ggml_context* ctx;
    bool ok = true;
    std::vector<void*> tmp;
    template <typename T>
    T* alloc(size_t n) {
        void* p = nullptr;
        if (cudaMalloc(&p, n * sizeof(T) + 64) != cudaSuccess) { cudaGetLastError(); ok = false; return nullptr; }
        tmp.push_back(p);
        return static_cast<T*>(p);
    }
    ~CondScratch() {
        for (void* p : tmp) cudaFree(p);
    }
};

// text encoder: token ids (+ optional key mask) -> fp32 hidden states [n, H] on the host (qwen_model.cpp:528-666)
ace_ggml_status text_forward_impl(ace_ggml_context* ctx, const int32_t* ids, const int32_t* mask, int n, int n_layers,
                                  bool apply_final_norm, bool embeddings_only, float* out, size_t out_size) {
    DeviceGuard guard(ctx->device);
    const QwenModel& q = ctx->text;
    const int H = q.cfg.hidden_size;
    const size_t needed = (size_t)H * n * sizeof(float);
    if (out_size < needed) return set_error(ctx, ACE_GGML_ERR_INVALID_ARG, "output buffer too small");
    for (int i = 0; i < n; ++i)
        if (ids[i] < 0 || ids[i] >= q.cfg.vocab_size) return set_error(ctx, ACE_GGML_ERR_INVALID_ARG, "token id out of range");
    std::string err;
    if (!enc_workspace_reserve(q.stack, ctx->ews, n, err)) return set_error(ctx, ACE_GGML_ERR, err);
    CondScratch sc{ctx};
    int32_t* d_ids = sc.alloc<int32_t>((size_t)n);
    float* d_out = sc.alloc<float>((size_t)n * H);
    if (!sc.ok) return set_error(ctx, ACE_GGML_ERR, "text encoder scratch allocation failed");
    cudaStream_t st = ctx->stream;
    cudaError_t e = cudaMemcpyAsync(d_ids, ids, (size_t)n * 4, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess && mask) e = cudaMemcpyAsync(ctx->ews.d_mask, mask, (size_t)n * 4, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return cuda_fail(ctx, "token upload failed", e);
    cudaEventRecord(ctx->ev0, st);
    launch_gather_rows(q.table, d_ids, ctx->ews.x, n, H, q.cfg.vocab_size, st);
    if (embeddings_only) {
        e = cudaMemcpyAsync(d_out, ctx->ews.x, needed, cudaMemcpyDeviceToDevice, st);
    } else {
        const int total = q.cfg.num_hidden_layers;
        const int run = (n_layers < 0 || n_layers > total) ? total : n_layers;
        // the final norm applies only when every layer ran (qwen_model.cpp:658-661)
        e = enc_stack_forward(q.stack, ctx->ews, n, run, /*causal=*/true, mask ? ctx->ews.d_mask : nullptr, apply_final_norm && run == total, d_out, st);
    }
    cudaEventRecord(ctx->ev1, st);
    if (e != cudaSuccess) return cuda_fail(ctx, "graph compute failed", e);
    e = cudaMemcpyAsync(out, d_out, needed, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return cuda_fail(ctx, "graph compute failed", e);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->last_device_ms = ms;
    return ACE_GGML_OK;
}

// y[n, N] = x[n, K] . W^T (+ bias): ace_project_tokens_linear (acestep_ggml.cpp:1625-1685), host in / host out
ace_ggml_status project_tokens(ace_ggml_context* ctx, const PackedWeight& W, const float* bias, const float* x, int n, int K,
                               std::vector<float>& out, int& out_dim) {
    if (W.K != K || W.N <= 0) return set_error(ctx, ACE_GGML_ERR, "linear projection weight shape mismatch");
    CondScratch sc{ctx};
    float* d_x = sc.alloc<float>((size_t)n * K);
    __half* d_x16 = sc.alloc<__half>((size_t)n * K);
    float* d_y = sc.alloc<float>((size_t)n * W.N);
    if (!sc.ok) return set_error(ctx, ACE_GGML_ERR, "failed to allocate projection input");
    cudaStream_t st = ctx->stream;
    cudaError_t e = cudaMemcpyAsync(d_x, x, (size_t)n * K * 4, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return cuda_fail(ctx, "projection upload failed", e);
    launch_f32_to_f16(d_x, d_x16, (long)n * K, st);
    GemmProblem g;
    g.A = d_x16;
    g.B = 1;
    g.L = n;
    g.C = K;
    g.a_lstride = K;
    g.a_bstride = (long)n * K;
    g.l_out = n;
    g.W = &W;
    g.epi.bias = bias;
    g.epi.out_f32 = d_y;
    g.epi.ld_out = W.N;
    g.epi.out_bstride = (long)n * W.N;
    e = gemm_launch(g, st);
    if (e != cudaSuccess) return cuda_fail(ctx, "linear projection compute failed", e);
    out_dim = W.N;
    out.resize((size_t)n * W.N);
    e = cudaMemcpyAsync(out.data(), d_y, out.size() * 4, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return cuda_fail(ctx, "linear projection compute failed", e);
    return ACE_GGML_OK;
}

// embed -> transformer stack -> final norm: forward_lyric_encoder / forward_timbre_encoder (acestep_dit_model.cpp:1562-1739).
// first_token_only: the timbre embedding is the first time step of the result (:1737-1738)
ace_ggml_status cond_encoder(ace_ggml_context* ctx, const EncStack& stack, const PackedWeight& proj, const float* proj_b, const float* x,
                             int n, int in_dim, const char* max_layers_env, bool first_token_only, std::vector<float>& out, int& out_dim) {
    const int H = ctx->dit.cfg.hidden_size;
    if (proj.K != in_dim || proj.N != H) return set_error(ctx, ACE_GGML_ERR, first_token_only ? "forward_timbre_encoder failed" : "forward_lyric_encoder failed");
    std::string err;
    if (!enc_workspace_reserve(stack, ctx->ews, n, err)) return set_error(ctx, ACE_GGML_ERR, err);
    CondScratch sc{ctx};
    float* d_x = sc.alloc<float>((size_t)n * in_dim);
    __half* d_x16 = sc.alloc<__half>((size_t)n * in_dim);
    float* d_y = sc.alloc<float>((size_t)n * H);
    if (!sc.ok) return set_error(ctx, ACE_GGML_ERR, "condition encoder scratch allocation failed");
    cudaStream_t st = ctx->stream;
    cudaError_t e = cudaMemcpyAsync(d_x, x, (size_t)n * in_dim * 4, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return cuda_fail(ctx, "condition upload failed", e);
    launch_f32_to_f16(d_x, d_x16, (long)n * in_dim, st);
    GemmProblem g;
    g.A = d_x16;
    g.B = 1;
    g.L = n;
    g.C = in_dim;
    g.a_lstride = in_dim;
    g.a_bstride = (long)n * in_dim;
    g.l_out = n;
    g.W = &proj;
    g.epi.bias = proj_b;
    g.epi.out_f32 = ctx->ews.x;
    g.epi.ld_out = H;
    g.epi.out_bstride = (long)n * H;
    e = gemm_launch(g, st);
    int max_layers = (int)stack.layers.size();
    if (const char* v = std::getenv(max_layers_env)) {
        char* endp = nullptr;
        const long long val = std::strtoll(v, &endp, 10);
        if (endp && endp != v && val >= 0) max_layers = (int)std::min<long long>(max_layers, val);
    }
    // the all-ones key mask the reference passes is a no-op; no mask, bidirectional, RoPE, sliding window per layer
    if (e == cudaSuccess) e = enc_stack_forward(stack, ctx->ews, n, max_layers, false, nullptr, true, d_y, st);
    if (e != cudaSuccess) return cuda_fail(ctx, first_token_only ? "timbre encoder compute failed" : "lyric encoder compute failed", e);
    const int rows = first_token_only ? 1 : n;
    out_dim = H;
    out.resize((size_t)rows * H);
    e = cudaMemcpyAsync(out.data(), d_y, out.size() * 4, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return cuda_fail(ctx, "condition encoder compute failed", e);
    return ACE_GGML_OK;
}

// [seq1 | seq2] with the valid tokens of both moved to the front, order preserved (acestep_ggml.cpp:1744-1801)
void pack_sequences(const std::vector<float>& h1, const std::vector<int32_t>& m1, int len1, const std::vector<float>& h2,
                    const std::vector<int32_t>& m2, int len2, int dim, std::vector<float>& out_h, std::vector<int32_t>& out_m) {
    if (len1 <= 0 && len2 <= 0) { out_h.clear(); out_m.clear(); return; }
    if (len2 <= 0) { out_h = h1; out_m = m1; return; }
    if (len1 <= 0) { out_h = h2; out_m = m2; return; }
    const int len = len1 + len2;
    out_h.assign((size_t)len * dim, 0.f);
    out_m.assign((size_t)len, 0);
    auto valid = [&](int i) { return i < len1 ? m1[(size_t)i] != 0 : m2[(size_t)(i - len1)] != 0; };
    int w = 0;
    for (int pass = 0; pass < 2; ++pass)      // pass 0: valid tokens in order; pass 1: the padded ones
        for (int i = 0; i < len; ++i) {
            if (valid(i) != (pass == 0)) continue;
            const float* src = i < len1 ? h1.data() + (size_t)i * dim : h2.data() + (size_t)(i - len1) * dim;
            std::memcpy(out_h.data() + (size_t)w * dim, src, (size_t)dim * 4);
            out_m[(size_t)w] = pass == 0 ? 1 : 0;
            ++w;
        }
}

// context_latents[T, ctx_dim] = [silence latent (src_dim) | chunk mask fill]  (acestep_ggml.cpp:1946-2040)
ace_ggml_status build_context_latents(ace_ggml_context* ctx, int seq_len, int audio_dim, int ctx_dim, int hop, int audio_channels,
                                      std::vector<float>& context) {
    context.assign((size_t)seq_len * ctx_dim, 0.f);
    if (read_nonneg_env_int("ACE_GGML_USE_SILENCE_CONTEXT", 1) == 0 || ctx_dim <= 0) return ACE_GGML_OK;
    const int src_dim = std::min(audio_dim, ctx_dim);
    std::vector<float> src((size_t)seq_len * src_dim, 0.f);
    const int debug_context = read_nonneg_env_int("ACE_GGML_DEBUG_CONTEXT", 0);
    const float fill = read_nonneg_env_int("ACE_GGML_CHUNK_MASK_FILL", 1) != 0 ? 1.0f : 0.0f;
    bool loaded = false;
    if (src_dim > 0) {
        if (const char* path = std::getenv("ACE_GGML_SILENCE_LATENT_F32")) {
            std::string err;
            if (load_silence_latent_f32(path, seq_len, src_dim, src, err)) {
                loaded = true;
                if (debug_context) std::fprintf(stderr, "ace_ggml: silence latent loaded from %s\n", path);
            } else if (debug_context) {
                std::fprintf(stderr, "ace_ggml: failed to load silence latent %s (%s)\n", path, err.c_str());
            }
        }
        if (!loaded) {
            // VAE-encode digital silence in chunks (default: 64 frames when seq_len > 128, else the whole clip).  Every chunk is the
            // same all-zero input, so each distinct chunk length is encoded once and reused.
            const int want = read_nonneg_env_int("ACE_GGML_VAE_ENCODE_CHUNK_FRAMES", 0);
            int chunk = seq_len;
            if (want > 0) chunk = std::max(1, std::min(want, seq_len));
            else {
                int a = read_nonneg_env_int("ACE_GGML_VAE_ENCODE_CHUNK_FRAMES_AUTO", 0);
                if (a <= 0) a = seq_len > 128 ? 64 : seq_len;
                chunk = std::max(1, std::min(a, seq_len));
            }
            bool ok = ctx->vae.has_encoder;
            std::vector<float> enc_full, enc_tail;
            auto encode_zeros = [&](int frames, std::vector<float>& dst) {
                std::vector<float> silence((size_t)frames * hop * audio_channels, 0.f);
                dst.assign((size_t)frames * audio_dim, 0.f);
                return vae_encode_impl(ctx, silence.data(), frames * hop, dst.data(), dst.size() * 4) == ACE_GGML_OK;
            };
            for (int f0 = 0; ok && f0 < seq_len; f0 += chunk) {
                const int cur = std::min(chunk, seq_len - f0);
                std::vector<float>& e = cur == chunk ? enc_full : enc_tail;
                if (e.empty()) ok = encode_zeros(cur, e);
                for (int t = 0; ok && t < cur; ++t)
                    std::memcpy(src.data() + (size_t)(f0 + t) * src_dim, e.data() + (size_t)t * audio_dim, (size_t)src_dim * 4);
            }
            if (ok) loaded = true;
            else {  // the reference keeps zero source latents when the encode fails (acestep_ggml.cpp:2019-2026)
                std::fill(src.begin(), src.end(), 0.f);
                ctx->last_error.clear();
                if (debug_context) std::fprintf(stderr, "ace_ggml: silence encode fallback failed, using zero src latents\n");
            }
        }
    }
    for (int t = 0; t < seq_len; ++t) {
        float* row = context.data() + (size_t)t * ctx_dim;
        if (src_dim > 0) std::memcpy(row, src.data() + (size_t)t * src_dim, (size_t)src_dim * 4);
        for (int c = src_dim; c < ctx_dim; ++c) row[c] = fill;
    }
    return ACE_GGML_OK;
}

double rms_of_frame(const float* xt, int seq_len, int dim, int t) {
    t = std::max(0, std::min(t, seq_len - 1));
    double acc = 0.0;
    for (int c = 0; c < dim; ++c) acc += (double)xt[(size_t)t * dim + c] * (double)xt[(size_t)t * dim + c];
    return std::sqrt(acc / (double)dim);
}

// noise -> 8 Euler steps -> VAE decode, everything between the two on the device (acestep_ggml.cpp:1901-2238)
ace_ggml_status generate_from_encoder(ace_ggml_context* ctx, const float* enc, const int32_t* enc_mask, int enc_len, int seq_len,
                                      float shift, int32_t seed, float* out_audio, size_t out_size, int32_t* out_samples,
                                      int32_t* out_channels) {
    if (!ctx || !enc || enc_len <= 0 || seq_len <= 0 || !out_audio) return ACE_GGML_ERR_INVALID_ARG;
    if (!ctx->dit_loaded) return set_error(ctx, ACE_GGML_ERR, "dit not loaded");
    if (!ctx->vae_loaded) return set_error(ctx, ACE_GGML_ERR, "vae not loaded");
    DeviceGuard guard(ctx->device);
    const int audio_dim = ctx->dit.cfg.audio_acoustic_hidden_dim;
    const int ctx_dim = ctx->dit.cfg.in_channels - audio_dim;
    if (ctx_dim <= 0) return set_error(ctx, ACE_GGML_ERR, "invalid dit context dimension");
    const VaeConfig& vc = ctx->vae.cfg;
    const int hop = vc.hop_length, ach = vc.audio_channels;
    if (vc.decoder_input_channels != audio_dim) return set_error(ctx, ACE_GGML_ERR, "vae latent channels mismatch with dit audio dim");
    const size_t needed = (size_t)seq_len * hop * ach * sizeof(float);
    if (out_size < needed) return set_error(ctx, ACE_GGML_ERR_INVALID_ARG, "output buffer too small");

    std::vector<float> context;
    ace_ggml_status s = build_context_latents(ctx, seq_len, audio_dim, ctx_dim, hop, ach, context);
    if (s != ACE_GGML_OK) return s;

    // seed-exact noise: the same generator / distribution objects the reference draws from, frame-major (acestep_ggml.cpp:2042-2047)
    std::mt19937 rng(static_cast<uint32_t>(seed));
    std::normal_distribution<float> dist(0.0f, 1.0f);
    std::vector<float> xt((size_t)seq_len * audio_dim);
    for (float& v : xt) v = dist(rng);

    std::vector<float> schedule;
    pick_shift_schedule(shift, schedule);
    std::vector<int32_t> attention_mask((size_t)seq_len, 1);
    s = sample_latents_device(ctx, xt.data(), context.data(), enc, nullptr, attention_mask.data(), enc_mask, 1, seq_len, enc_len,
                              schedule.data(), (int)schedule.size(), 1.0f);
    if (s != ACE_GGML_OK) return s;

    const int debug_stats = read_nonneg_env_int("ACE_GGML_DEBUG_GENERATE_STATS", 0);
    if (debug_stats) {
        cudaError_t e = cudaMemcpyAsync(xt.data(), ctx->d_x, xt.size() * 4, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) return cuda_fail(ctx, "latent download failed", e);
        std::fprintf(stderr, "ace_ggml: xt_rms t0=%.6f t25%%=%.6f t50%%=%.6f t75%%=%.6f tEnd=%.6f\n", rms_of_frame(xt.data(), seq_len, audio_dim, 0),
                     rms_of_frame(xt.data(), seq_len, audio_dim, seq_len / 4), rms_of_frame(xt.data(), seq_len, audio_dim, seq_len / 2),
                     rms_of_frame(xt.data(), seq_len, audio_dim, (seq_len * 3) / 4), rms_of_frame(xt.data(), seq_len, audio_dim, seq_len - 1));
        std::fflush(stderr);
    }

    // VAE decode.  The reference tiles clips longer than 128 frames (chunk 128 / overlap 32, acestep_ggml.cpp:2114-2229); its halo
    // exceeds the decoder's receptive field (~9.5 frames), so the tiled result equals an un-tiled decode, which is what runs here.
    // A caller that SETS ACE_GGML_VAE_CHUNK_FRAMES / _OVERLAP_FRAMES gets exactly that window policy (a short overlap changes the
    // audio near the seams, and a drop-in must reproduce it).
    if (!ctx->io.reserve(0, needed)) return set_error(ctx, ACE_GGML_ERR, "audio buffer allocation failed");
    float* d_audio = static_cast<float*>(ctx->io.d_out);
    const char* chunk_env = std::getenv("ACE_GGML_VAE_CHUNK_FRAMES");
    const char* overlap_env = std::getenv("ACE_GGML_VAE_CHUNK_OVERLAP_FRAMES");
    int chunk = 0;
    if (chunk_env && chunk_env[0]) chunk = std::max(0, read_nonneg_env_int("ACE_GGML_VAE_CHUNK_FRAMES", 0));
    else if (overlap_env && overlap_env[0]) {
        const int a = read_nonneg_env_int("ACE_GGML_VAE_CHUNK_FRAMES_AUTO", 0);
        chunk = a > 0 ? a : (seq_len > 128 ? 128 : 0);
    }
    std::string err;
    cudaError_t e = cudaSuccess;
    long produced = (long)seq_len * hop;
    if (chunk > 0 && chunk < seq_len) {
        int overlap = read_nonneg_env_int("ACE_GGML_VAE_CHUNK_OVERLAP_FRAMES", -1);
        if (overlap < 0) overlap = std::min(64, std::max(1, chunk / 4));
        if (overlap * 2 >= chunk) overlap = std::max(0, chunk / 2 - 1);
        int stride = chunk - 2 * overlap;
        if (stride <= 0) { overlap = 0; stride = chunk; }
        const size_t win_audio = (size_t)(chunk + 2 * overlap + stride) * hop * ach;
        CondScratch sc{ctx};
        float* d_win = sc.alloc<float>(win_audio);
        if (!sc.ok) return set_error(ctx, ACE_GGML_ERR, "audio buffer allocation failed");
        size_t write_pos = 0;
        int idx = 0;
        for (int core0 = 0; core0 < seq_len && e == cudaSuccess; core0 += stride, ++idx) {
            const int core1 = std::min(core0 + stride, seq_len);
            const int w0 = std::max(0, core0 - overlap), w1 = std::min(seq_len, core1 + overlap);
            e = vae_decode_grouped(ctx, ctx->d_x + (size_t)w0 * audio_dim, 1, w1 - w0, d_win, err);
            if (e != cudaSuccess) break;
            const size_t trim0 = (size_t)(core0 - w0) * hop, core = (size_t)(core1 - core0) * hop;  // upsample factor is exactly hop
            e = cudaMemcpyAsync(d_audio + write_pos * ach, d_win + trim0 * ach, core * ach * 4, cudaMemcpyDeviceToDevice, ctx->stream);
            write_pos += core;
        }
        produced = (long)write_pos;
    } else {
        e = vae_decode_grouped(ctx, ctx->d_x, 1, seq_len, d_audio, err);
    }
    cudaEventRecord(ctx->ev1, ctx->stream);
    if (e != cudaSuccess) return err.empty() ? cuda_fail(ctx, "vae decode launch failed", e) : set_error(ctx, ACE_GGML_ERR, err);
    e = cudaMemcpyAsync(out_audio, d_audio, (size_t)produced * ach * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) return cuda_fail(ctx, "generate_audio failed", e);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->last_device_ms = ms;
    if (debug_stats) {
        double acc = 0.0, peak = 0.0;
        const size_t n = (size_t)produced * ach;
        for (size_t i = 0; i < n; ++i) {
            acc += (double)out_audio[i] * (double)out_audio[i];
            peak = std::max(peak, std::fabs((double)out_audio[i]));
        }
        std::fprintf(stderr, "ace_ggml: vae_chunk idx=0 frame=[0,%d) rms=%.6f peak=%.6f\n", seq_len, n ? std::sqrt(acc / (double)n) : 0.0, peak);
        std::fprintf(stderr, "ace_ggml: vae_chunk_concat samples=%ld expected=%d factor=%.6f\n", produced, seq_len * hop, (double)hop);
        std::fflush(stderr);
    }
    if (out_samples) *out_samples = (int32_t)produced;
    if (out_channels) *out_channels = ach;
    return ACE_GGML_OK;
}

// text states of one prompt: all layers, or ACE_GGML_TEXT_MAX_LAYERS of them (acestep_ggml.cpp:2274-2297)
ace_ggml_status style_text_states(ace_ggml_context* ctx, const int32_t* ids, int n, std::vector<float>& out) {
    const int H = ctx->text.cfg.hidden_size;
    out.assign((size_t)n * H, 0.f);
    const int max_layers = read_nonneg_env_int("ACE_GGML_TEXT_MAX_LAYERS", -1);
    if (max_layers >= 0) return text_forward_impl(ctx, ids, nullptr, n, max_layers, true, false, out.data(), out.size() * 4);
    return text_forward_impl(ctx, ids, nullptr, n, -1, true, false, out.data(), out.size() * 4);
}

ace_ggml_status generate_style_lyric_timbre(ace_ggml_context* ctx, const int32_t* style_ids, int n_style, const int32_t* lyric_ids, int n_lyric,
                                            const float* refer, const int32_t* refer_order_mask, int n_refer, int refer_len, int seq_len,
                                            float shift, int32_t seed, float* out_audio, size_t out_size, int32_t* out_samples,
                                            int32_t* out_channels) {
    const bool has_style = n_style > 0, has_lyric = n_lyric > 0, has_timbre = n_refer > 0;
    if (!ctx || !out_audio || seq_len <= 0 || (!has_style && !has_lyric && !has_timbre)) return ACE_GGML_ERR_INVALID_ARG;
    if ((has_style && !style_ids) || (has_lyric && !lyric_ids) || (has_timbre && (!refer || refer_len <= 0))) return ACE_GGML_ERR_INVALID_ARG;
    if (!ctx->text.loaded) return set_error(ctx, ACE_GGML_ERR, "text encoder not loaded");
    if (!ctx->dit_loaded) return set_error(ctx, ACE_GGML_ERR, "dit not loaded");
    if (!ctx->vae_loaded) return set_error(ctx, ACE_GGML_ERR, "vae not loaded");
    DeviceGuard guard(ctx->device);
    const DitModel& dit = ctx->dit;
    const int text_hidden = ctx->text.cfg.hidden_size, H = dit.cfg.hidden_size;
    const bool allow_mismatch = std::getenv("ACE_GGML_ALLOW_TEXT_DIM_MISMATCH") != nullptr;
    ace_ggml_status st = ACE_GGML_OK;

    std::vector<float> style_states, lyric_states;
    if (has_style) {
        st = style_text_states(ctx, style_ids, n_style, style_states);
        if (st != ACE_GGML_OK) return st;
    }
    if (has_lyric) {  // lyrics enter as raw token embeddings (acestep_ggml.cpp:2398-2409)
        lyric_states.assign((size_t)n_lyric * text_hidden, 0.f);
        st = text_forward_impl(ctx, lyric_ids, nullptr, n_lyric, 0, false, true, lyric_states.data(), lyric_states.size() * 4);
        if (st != ACE_GGML_OK) return st;
    }
    std::vector<float> style_cond, lyric_cond, timbre_cond;
    std::vector<int32_t> timbre_mask;
    int style_dim = 0, lyric_dim = 0, timbre_dim = 0;
    bool style_encoded = false, lyric_encoded = false;
    if (has_style && dit.has_text_projector) {
        st = project_tokens(ctx, dit.text_projector, nullptr, style_states.data(), n_style, text_hidden, style_cond, style_dim);
        if (st != ACE_GGML_OK) return st;
        style_encoded = style_dim == H;
    }
    if (has_lyric && (dit.has_lyric_embed || dit.has_text_projector)) {
        const int in_dim = dit.cfg.text_hidden_dim > 0 ? dit.cfg.text_hidden_dim : 1024;
        const PackedWeight& pw = dit.has_lyric_embed ? dit.lyric_embed : dit.text_projector;
        if (in_dim == text_hidden)
            st = cond_encoder(ctx, dit.lyric, pw, dit.has_lyric_embed ? dit.lyric_embed_b : nullptr, lyric_states.data(), n_lyric, in_dim,
                              "ACE_GGML_LYRIC_MAX_LAYERS", false, lyric_cond, lyric_dim);
        else st = set_error(ctx, ACE_GGML_ERR, "forward_lyric_encoder failed");
        if (st == ACE_GGML_OK && lyric_dim == H) lyric_encoded = true;
        else if (st != ACE_GGML_OK) ctx->last_error.clear();  // the reference falls back to the raw states (:2449-2452)
    }
    if (has_timbre) {
        if (!dit.has_timbre_embed) return set_error(ctx, ACE_GGML_ERR, "timbre encoder weights are not loaded");
        const int in_dim = dit.cfg.timbre_hidden_dim > 0 ? dit.cfg.timbre_hidden_dim : (dit.cfg.audio_acoustic_hidden_dim > 0 ? dit.cfg.audio_acoustic_hidden_dim : 64);
        for (int i = 0; i < n_refer; ++i) {
            if (refer_order_mask && refer_order_mask[i] != 0)
                return set_error(ctx, ACE_GGML_ERR_UNSUPPORTED, "multi-batch refer_audio_order_mask is not supported");
            std::vector<float> one;
            int dim = 0;
            st = cond_encoder(ctx, dit.timbre, dit.timbre_embed, dit.timbre_embed_b, refer + (size_t)i * refer_len * in_dim, refer_len, in_dim,
                              "ACE_GGML_TIMBRE_MAX_LAYERS", true, one, dim);
            if (st != ACE_GGML_OK) return st;
            timbre_dim = dim;
            timbre_cond.insert(timbre_cond.end(), one.begin(), one.end());
            timbre_mask.push_back(1);
        }
        if (timbre_dim != H) return set_error(ctx, ACE_GGML_ERR, "timbre condition dim mismatch with dit");
    }
    if ((has_style && !style_encoded) || (has_lyric && !lyric_encoded)) {
        if (text_hidden != H && !allow_mismatch) return set_error(ctx, ACE_GGML_ERR, "text encoder hidden size mismatch with dit");
    }
    const int cpy = std::min(text_hidden, H);
    auto widen = [&](const std::vector<float>& states, const std::vector<float>& cond, bool encoded, int n, std::vector<float>& dst) {
        dst.assign((size_t)n * H, 0.f);
        if (encoded) std::memcpy(dst.data(), cond.data(), cond.size() * 4);
        else
            for (int t = 0; t < n; ++t) std::memcpy(dst.data() + (size_t)t * H, states.data() + (size_t)t * text_hidden, (size_t)cpy * 4);
    };
    std::vector<float> enc, style_hidden;
    std::vector<int32_t> enc_mask, style_mask;
    int cond_len = 0;
    if (has_lyric) {
        widen(lyric_states, lyric_cond, lyric_encoded, n_lyric, enc);
        enc_mask.assign((size_t)n_lyric, 1);
        cond_len = n_lyric;
    }
    if (has_timbre) {  // order: lyric, timbre, style (acestep_ggml.cpp:2505-2551)
        std::vector<float> ph;
        std::vector<int32_t> pm;
        pack_sequences(enc, enc_mask, cond_len, timbre_cond, timbre_mask, (int)timbre_mask.size(), H, ph, pm);
        enc.swap(ph);
        enc_mask.swap(pm);
        cond_len += (int)timbre_mask.size();
    }
    if (has_style) {
        widen(style_states, style_cond, style_encoded, n_style, style_hidden);
        style_mask.assign((size_t)n_style, 1);
        std::vector<float> ph;
        std::vector<int32_t> pm;
        pack_sequences(enc, enc_mask, cond_len, style_hidden, style_mask, n_style, H, ph, pm);
        enc.swap(ph);
        enc_mask.swap(pm);
        cond_len += n_style;
    }
    if (cond_len <= 0) return set_error(ctx, ACE_GGML_ERR_INVALID_ARG, "empty style/lyric/timbre inputs");
    return generate_from_encoder(ctx, enc.data(), enc_mask.data(), cond_len, seq_len, shift, seed, out_audio, out_size, out_samples, out_channels);
}

}  // namespace

extern "C" {

ace_ggml_status ace_ggml_load_lm(ace_ggml_context* ctx, const char* model_dir) {
    if (!ctx || !model_dir) return ACE_GGML_ERR_INVALID_ARG;
    DeviceGuard guard(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    std::string err;
    const int st = qwen_load(model_dir, ctx->text, err);  // the reference loads into the text-encoder slot too (acestep_ggml.cpp:246-251)
    if (st != 0) return set_error(ctx, map_load_status(st), err);
    return ACE_GGML_OK;
}

ace_ggml_status ace_ggml_load_text_encoder(ace_ggml_context* ctx, const char* model_dir) {
    if (!ctx || !model_dir) return ACE_GGML_ERR_INVALID_ARG;
    DeviceGuard guard(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    std::string err;
    const int st = qwen_load(model_dir, ctx->text, err);
    if (st != 0) return set_error(ctx, map_load_status(st), err);
    return ACE_GGML_OK;
}

ace_ggml_status ace_ggml_vae_encode(ace_ggml_context* ctx, const float* audio, int32_t n_samples, float* out, size_t out_size) {
    if (!ctx || !audio || !out || n_samples <= 0) return ACE_GGML_ERR_INVALID_ARG;
    if (!ctx->vae_loaded) return set_error(ctx, ACE_GGML_ERR, "vae not loaded");
    return vae_encode_impl(ctx, audio, n_samples, out, out_size);
}

ace_ggml_status ace_ggml_text_encoder_forward(ace_ggml_context* ctx, const int32_t* token_ids, int32_t n_tokens, float* out, size_t out_size) {
    if (!ctx || !token_ids || !out || n_tokens <= 0) return ACE_GGML_ERR_INVALID_ARG;
    if (!ctx->text.loaded) return set_error(ctx, ACE_GGML_ERR, "text encoder not loaded");
    return text_forward_impl(ctx, token_ids, nullptr, n_tokens, -1, true, false, out, out_size);
}
ace_ggml_status ace_ggml_text_encoder_forward_masked(ace_ggml_context* ctx, const int32_t* token_ids, const int32_t* attention_mask,
                                                     int32_t n_tokens, float* out, size_t out_size) {
    if (!ctx || !token_ids || !out || n_tokens <= 0) return ACE_GGML_ERR_INVALID_ARG;
    if (!ctx->text.loaded) return set_error(ctx, ACE_GGML_ERR, "text encoder not loaded");
    return text_forward_impl(ctx, token_ids, attention_mask, n_tokens, -1, true, false, out, out_size);
}
ace_ggml_status ace_ggml_text_encoder_forward_embeddings(ace_ggml_context* ctx, const int32_t* token_ids, int32_t n_tokens, float* out,
                                                         size_t out_size) {
    if (!ctx || !token_ids || !out || n_tokens <= 0) return ACE_GGML_ERR_INVALID_ARG;
    if (!ctx->text.loaded) return set_error(ctx, ACE_GGML_ERR, "text encoder not loaded");
    return text_forward_impl(ctx, token_ids, nullptr, n_tokens, 0, false, true, out, out_size);
}
ace_ggml_status ace_ggml_text_encoder_forward_layers(ace_ggml_context* ctx, const int32_t* token_ids, const int32_t* attention_mask,
                                                     int32_t n_tokens, int32_t n_layers, int32_t apply_final_norm, float* out, size_t out_size) {
    if (!ctx || !token_ids || !out || n_tokens <= 0) return ACE_GGML_ERR_INVALID_ARG;
    if (!ctx->text.loaded) return set_error(ctx, ACE_GGML_ERR, "text encoder not loaded");
    return text_forward_impl(ctx, token_ids, attention_mask, n_tokens, n_layers, apply_final_norm != 0, false, out, out_size);
}

ace_ggml_status ace_ggml_generate_audio_simple(ace_ggml_context* ctx, const int32_t* token_ids, int32_t n_tokens, int32_t seq_len, float shift,
                                               int32_t seed, float* out_audio, size_t out_size, int32_t* out_audio_samples,
                                               int32_t* out_audio_channels) {
    if (!ctx || !token_ids || !out_audio || n_tokens <= 0 || seq_len <= 0) return ACE_GGML_ERR_INVALID_ARG;
    if (!ctx->text.loaded) return set_error(ctx, ACE_GGML_ERR, "text encoder not loaded");
    if (!ctx->dit_loaded) return set_error(ctx, ACE_GGML_ERR, "dit not loaded");
    if (!ctx->vae_loaded) return set_error(ctx, ACE_GGML_ERR, "vae not loaded");
    const int H = ctx->dit.cfg.hidden_size, text_hidden = ctx->text.cfg.hidden_size;
    if (text_hidden != H && !std::getenv("ACE_GGML_ALLOW_TEXT_DIM_MISMATCH"))
        return set_error(ctx, ACE_GGML_ERR, "text encoder hidden size mismatch with dit");
    std::vector<float> text_states;
    ace_ggml_status st = style_text_states(ctx, token_ids, n_tokens, text_states);
    if (st != ACE_GGML_OK) return st;
    std::vector<float> enc((size_t)n_tokens * H, 0.f);
    const int cpy = std::min(text_hidden, H);
    for (int t = 0; t < n_tokens; ++t) std::memcpy(enc.data() + (size_t)t * H, text_states.data() + (size_t)t * text_hidden, (size_t)cpy * 4);
    std::vector<int32_t> mask((size_t)n_tokens, 1);
    return generate_from_encoder(ctx, enc.data(), mask.data(), n_tokens, seq_len, shift, seed, out_audio, out_size, out_audio_samples, out_audio_channels);
}

ace_ggml_status ace_ggml_generate_audio_style_lyric_simple(ace_ggml_context* ctx, const int32_t* style_token_ids, int32_t n_style_tokens,
                                                           const int32_t* lyric_token_ids, int32_t n_lyric_tokens, int32_t seq_len, float shift,
                                                           int32_t seed, float* out_audio, size_t out_size, int32_t* out_audio_samples,
                                                           int32_t* out_audio_channels) {
    return generate_style_lyric_timbre(ctx, style_token_ids, n_style_tokens, lyric_token_ids, n_lyric_tokens, nullptr, nullptr, 0, 0, seq_len, shift,
                                       seed, out_audio, out_size, out_audio_samples, out_audio_channels);
}

ace_ggml_status ace_ggml_generate_audio_style_lyric_timbre_simple(ace_ggml_context* ctx, const int32_t* style_token_ids, int32_t n_style_tokens,
                                                                  const int32_t* lyric_token_ids, int32_t n_lyric_tokens,
                                                                  const float* refer_audio_acoustic_hidden_states,
                                                                  const int32_t* refer_audio_order_mask, int32_t n_refer_audio,
                                                                  int32_t refer_audio_len, int32_t seq_len, float shift, int32_t seed,
                                                                  float* out_audio, size_t out_size, int32_t* out_audio_samples,
                                                                  int32_t* out_audio_channels) {
    return generate_style_lyric_timbre(ctx, style_token_ids, n_style_tokens, lyric_token_ids, n_lyric_tokens, refer_audio_acoustic_hidden_states,
                                       refer_audio_order_mask, n_refer_audio, refer_audio_len, seq_len, shift, seed, out_audio, out_size,
                                       out_audio_samples, out_audio_channels);
}

// additive: the sampler for callers that already hold encoder states (what ace_ggml_generate_audio_* run after their encoders)
ace_ggml_status ace_b200_generate_audio_from_encoder(ace_ggml_context* ctx, const float* encoder_hidden_states, const int32_t* encoder_attention_mask,
                                                     int32_t enc_len, int32_t seq_len, float shift, int32_t seed, float* out_audio, size_t out_size,
                                                     int32_t* out_audio_samples, int32_t* out_audio_channels) {
    return generate_from_encoder(ctx, encoder_hidden_states, encoder_attention_mask, enc_len, seq_len, shift, seed, out_audio, out_size,
                                 out_audio_samples, out_audio_channels);
}

// additive: the two condition encoders on their own (unit tests / callers that cache conditioning)
ace_ggml_status ace_b200_lyric_encoder_forward(ace_ggml_context* ctx, const float* lyric_embeds, int32_t n_tokens, float* out, size_t out_size) {
    if (!ctx || !lyric_embeds || !out || n_tokens <= 0) return ACE_GGML_ERR_INVALID_ARG;
    if (!ctx->dit_loaded) return set_error(ctx, ACE_GGML_ERR, "dit not loaded");
    DeviceGuard guard(ctx->device);
    const DitModel& dit = ctx->dit;
    if (!dit.has_lyric_embed && !dit.has_text_projector) return set_error(ctx, ACE_GGML_ERR, "forward_lyric_encoder failed");
    const int in_dim = dit.cfg.text_hidden_dim > 0 ? dit.cfg.text_hidden_dim : 1024;
    std::vector<float> y;
    int dim = 0;
    ace_ggml_status st = cond_encoder(ctx, dit.lyric, dit.has_lyric_embed ? dit.lyric_embed : dit.text_projector, dit.has_lyric_embed ? dit.lyric_embed_b : nullptr,
                                      lyric_embeds, n_tokens, in_dim, "ACE_GGML_LYRIC_MAX_LAYERS", false, y, dim);
    if (st != ACE_GGML_OK) return st;
    if (out_size < y.size() * 4) return set_error(ctx, ACE_GGML_ERR_INVALID_ARG, "output buffer too small");
    std::memcpy(out, y.data(), y.size() * 4);
    return ACE_GGML_OK;
}
ace_ggml_status ace_b200_timbre_encoder_forward(ace_ggml_context* ctx, const float* refer_states, int32_t refer_len, float* out, size_t out_size) {
    if (!ctx || !refer_states || !out || refer_len <= 0) return ACE_GGML_ERR_INVALID_ARG;
    if (!ctx->dit_loaded) return set_error(ctx, ACE_GGML_ERR, "dit not loaded");
    DeviceGuard guard(ctx->device);
    const DitModel& dit = ctx->dit;
    if (!dit.has_timbre_embed) return set_error(ctx, ACE_GGML_ERR, "timbre encoder weights are not loaded");
    const int in_dim = dit.cfg.timbre_hidden_dim > 0 ? dit.cfg.timbre_hidden_dim : (dit.cfg.audio_acoustic_hidden_dim > 0 ? dit.cfg.audio_acoustic_hidden_dim : 64);
    std::vector<float> y;
    int dim = 0;
    ace_ggml_status st = cond_encoder(ctx, dit.timbre, dit.timbre_embed, dit.timbre_embed_b, refer_states, refer_len, in_dim, "ACE_GGML_TIMBRE_MAX_LAYERS", true, y, dim);
    if (st != ACE_GGML_OK) return st;
    if (out_size < y.size() * 4) return set_error(ctx, ACE_GGML_ERR_INVALID_ARG, "output buffer too small");
    std::memcpy(out, y.data(), y.size() * 4);
    return ACE_GGML_OK;
}

}  // extern "C"
