// C ABI of libacestep_b200.so: the reference's ace_ggml_* entry points (include/acestep_ggml.h) plus the
// ace_b200_* extensions (include/acestep_b200.h).  Host pointers in, host pointers out; no exception crosses.
// reference behaviour mirrored: acestep_ggml.cpp:65-70, 192-236 (context, errors), 260-357 / 529-632 (load),
// 634-657 (get_info), 659-682 (vae_decode checks), 1304-1329 (dit_forward checks), 2056-2086 (Euler loop)
#include "abi_internal.h"

#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

namespace ace {
std::atomic<uint64_t> g_kernel_launches{0};
}

namespace ace_abi {

// memcmp == 0 over large host buffers on several threads (a 67 MB batch of encoder states takes ~6 ms on one core, longer than
// uploading it): 4 MB pieces dealt round-robin, every thread stops at the first difference anyone finds
bool host_bytes_equal(const void* a, const void* b, size_t n) {
    constexpr size_t PIECE = 4u << 20;
    if (n < 2 * PIECE) return std::memcmp(a, b, n) == 0;
    const size_t pieces = (n + PIECE - 1) / PIECE;
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const unsigned nt = (unsigned)std::min<size_t>(std::min(8u, hw), pieces);
    std::atomic<bool> differ{false};
    auto work = [&](unsigned t) {
        for (size_t i = t; i < pieces && !differ.load(std::memory_order_relaxed); i += nt) {
            const size_t off = i * PIECE, len = std::min(PIECE, n - off);
            if (std::memcmp(static_cast<const char*>(a) + off, static_cast<const char*>(b) + off, len) != 0) differ.store(true, std::memory_order_relaxed);
        }
    };
    std::vector<std::thread> pool;
    try {
        for (unsigned t = 1; t < nt; ++t) pool.emplace_back(work, t);
    } catch (...) {  // no threads to be had: the calling thread covers what was not started
        for (auto& th : pool) th.join();
        return std::memcmp(a, b, n) == 0;
    }
    work(0);
    for (auto& th : pool) th.join();
    return !differ.load();
}

ace_ggml_status stage_dit_conditioning(ace_ggml_context* ctx, const float* enc, const int32_t* enc_mask, int B, int E);

// H2D copies of one batched forward's inputs into the workspace
ace_ggml_status stage_dit_inputs(ace_ggml_context* ctx, const float* hidden, const float* context, const float* enc,
                                 const int32_t* mask, const int32_t* enc_mask, int B, int T, int E) {
    const DitConfig& c = ctx->dit.cfg;
    const int A = c.audio_acoustic_hidden_dim, cd = c.in_channels - A;
    DitWorkspace& ws = ctx->dws;
    cudaStream_t st = ctx->stream;
    cudaError_t e = cudaSuccess;
    auto h2d = [&](void* dst, const void* src, size_t bytes) {
        if (e == cudaSuccess && bytes) e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st);
    };
    if (hidden) h2d(ws.d_hidden, hidden, (size_t)B * T * A * 4);
    else if (e == cudaSuccess) e = cudaMemsetAsync(ws.d_hidden, 0, (size_t)B * T * A * 4, st);
    if (context) h2d(ws.d_context, context, (size_t)B * T * cd * 4);
    else if (e == cudaSuccess) e = cudaMemsetAsync(ws.d_context, 0, (size_t)B * T * cd * 4, st);
    if (mask) h2d(ws.d_mask, mask, (size_t)B * T * 4);
    if (e != cudaSuccess) return cuda_fail(ctx, "input upload failed", e);
    return E > 0 ? stage_dit_conditioning(ctx, enc, enc_mask, B, E) : ACE_GGML_OK;
}

// the step-invariant part of the inputs: encoder states (zeros when absent) and their mask
ace_ggml_status stage_dit_conditioning(ace_ggml_context* ctx, const float* enc, const int32_t* enc_mask, int B, int E) {
    const size_t n = (size_t)B * E * ctx->dit.cfg.hidden_size * 4;
    DitWorkspace& ws = ctx->dws;
    cudaError_t e = enc ? cudaMemcpyAsync(ws.d_enc, enc, n, cudaMemcpyHostToDevice, ctx->stream) : cudaMemsetAsync(ws.d_enc, 0, n, ctx->stream);
    if (e == cudaSuccess && enc_mask) e = cudaMemcpyAsync(ws.d_enc_mask, enc_mask, (size_t)B * E * 4, cudaMemcpyHostToDevice, ctx->stream);
    if (e != cudaSuccess) return cuda_fail(ctx, "input upload failed", e);
    return ACE_GGML_OK;
}

int env_max_layers() {
    const char* v = std::getenv("ACE_GGML_DIT_MAX_LAYERS");
    if (!v) return 0;
    char* endp = nullptr;
    const long long x = std::strtoll(v, &endp, 10);
    return (endp && endp != v && x > 0) ? (int)std::min<long long>(x, 1 << 20) : 0;
}

ace_ggml_status dit_forward_impl(ace_ggml_context* ctx, const float* hidden, const float* context, const float* enc,
                                 const int32_t* mask, const int32_t* enc_mask, int B, int T, int E, const float* t,
                                 const float* t_r, float* out, size_t out_size, const char* profile_name) {
    const auto t0 = Clock::now();
    DeviceGuard guard(ctx->device);
    const DitConfig& c = ctx->dit.cfg;
    const size_t needed = (size_t)B * T * c.audio_acoustic_hidden_dim * sizeof(float);
    if (out_size < needed) return set_error(ctx, ACE_GGML_ERR_INVALID_ARG, "output buffer too small");
    std::string err;
    if (!dit_workspace_reserve(ctx->dit, ctx->dws, B, T, E, err)) return set_error(ctx, ACE_GGML_ERR, err);
    const auto t_init = Clock::now();
    // step-invariant conditioning: same bytes as the previous call -> keep the device copies and the cached cross K/V
    DitForwardOpts o;
    o.max_layers = env_max_layers();
    const size_t enc_n = (size_t)B * E * c.hidden_size;
    bool same_cond = E > 0 && enc != nullptr && ctx->dws.cross_kv_valid && ctx->enc_cache_B == B && ctx->enc_cache_E == E &&
                     ctx->enc_cache_layers == o.max_layers && ctx->enc_cache.size() == enc_n && ctx->enc_cache_has_mask == (enc_mask != nullptr) &&
                     !env_on("ACE_B200_NO_COND_CACHE");
    // the per-step inputs go first: their (asynchronous, when the caller's memory is pinned) upload runs under the host-side compare
    ace_ggml_status s = stage_dit_inputs(ctx, hidden, context, nullptr, mask, nullptr, B, T, 0);
    if (s != ACE_GGML_OK) return s;
    if (same_cond) same_cond = host_bytes_equal(ctx->enc_cache.data(), enc, enc_n * 4);
    if (same_cond && enc_mask) same_cond = host_bytes_equal(ctx->enc_mask_cache.data(), enc_mask, (size_t)B * E * 4);
    if (!same_cond) {
        ctx->dws.cross_kv_valid = false;
        ctx->enc_cache_B = 0;
        if (E > 0 && enc != nullptr && !env_on("ACE_B200_NO_COND_CACHE")) {
            ctx->enc_cache.assign(enc, enc + enc_n);
            if (enc_mask) ctx->enc_mask_cache.assign(enc_mask, enc_mask + (size_t)B * E);
            ctx->enc_cache_has_mask = enc_mask != nullptr;
            ctx->enc_cache_B = B;
            ctx->enc_cache_E = E;
            ctx->enc_cache_layers = o.max_layers;
        }
    }
    if (!same_cond && E > 0) {
        s = stage_dit_conditioning(ctx, enc, enc_mask, B, E);
        if (s != ACE_GGML_OK) return s;
    }
    std::vector<float> tt((size_t)2 * B);
    for (int b = 0; b < B; ++b) {
        tt[(size_t)b] = t[b];
        tt[(size_t)B + b] = t[b] - t_r[b];  // second embedder sees t - t_r (acestep_dit_model.cpp:1423)
    }
    cudaError_t e = cudaMemcpyAsync(ctx->dws.d_t, tt.data(), tt.size() * 4, cudaMemcpyHostToDevice, ctx->stream);
    if (e != cudaSuccess) return cuda_fail(ctx, "timestep upload failed", e);
    o.has_mask = mask != nullptr;
    o.has_enc_mask = enc_mask != nullptr && E > 0;
    o.reuse_cross_kv = same_cond;
    cudaEventRecord(ctx->ev0, ctx->stream);
    e = dit_forward_device(ctx->dit, ctx->dws, B, T, E, ctx->dws.d_out, o, ctx->stream);
    cudaEventRecord(ctx->ev1, ctx->stream);
    if (e != cudaSuccess) return cuda_fail(ctx, "dit forward launch failed", e);
    // the D2H rides the same stream (no second blocking copy): truly asynchronous into pinned caller memory, staged by the
    // driver for pageable memory; one synchronise ends the call (and makes the pageable `tt` upload safe to drop)
    e = cudaMemcpyAsync(out, ctx->dws.d_out, needed, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) return cuda_fail(ctx, "dit forward failed", e);
    const auto t_done = Clock::now();
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->last_device_ms = ms;
    if (env_on("ACE_GGML_DIT_PROFILE")) {
        const auto t_end = Clock::now();
        std::fprintf(stderr,
                     "%s profile: backend=%s seq=%d enc=%d threads=%d init_ms=%lld build_ms=%lld graph_ms=%lld compute_ms=%lld "
                     "copy_ms=%lld total_ms=%lld device_ms=%.3f batch=%d\n",
                     profile_name, "b200", T, E, ctx->n_threads, ms_between(t0, t_init), 0LL, 0LL, ms_between(t_init, t_done),
                     ms_between(t_done, t_end), ms_between(t0, t_end), ms, B);
    }
    return ACE_GGML_OK;
}

// Clips of one call are decoded in groups whose activation workspace stays under ~40 GB (a 240 s clip needs ~13 GB of
// fp32 / fp16 level buffers; every decode already fills the machine, so grouping costs nothing).  ACE_B200_VAE_GROUP_FRAMES
// overrides the cap (frames per group, summed over the clips of the group).
int vae_group_size(int B, int n_frames) {
    long cap = 18000;
    if (const char* v = std::getenv("ACE_B200_VAE_GROUP_FRAMES")) cap = std::max(1L, std::atol(v));
    return (int)std::max(1L, std::min<long>(B, cap / std::max(1, n_frames)));
}

// latents: device [B, n_frames, C] -> audio: device [B, n_frames*hop, ch]; both caller-owned device buffers
cudaError_t vae_decode_grouped(ace_ggml_context* ctx, const float* d_latents, int B, int n_frames, float* d_audio, std::string& err) {
    const VaeConfig& c = ctx->vae.cfg;
    const int G = vae_group_size(B, n_frames);
    if (!vae_workspace_reserve(ctx->vae, ctx->vws, G, n_frames, err)) return cudaErrorMemoryAllocation;
    const size_t lat_per = (size_t)n_frames * c.decoder_input_channels, aud_per = (size_t)n_frames * c.hop_length * c.audio_channels;
    for (int b0 = 0; b0 < B; b0 += G) {
        const int g = std::min(G, B - b0);
        cudaError_t e = cudaMemcpyAsync(ctx->vws.d_lat, d_latents + (size_t)b0 * lat_per, (size_t)g * lat_per * 4, cudaMemcpyDeviceToDevice, ctx->stream);
        if (e == cudaSuccess) e = vae_decode_device(ctx->vae, ctx->vws, g, n_frames, d_audio + (size_t)b0 * aud_per, ctx->stream);
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

ace_ggml_status vae_decode_impl(ace_ggml_context* ctx, const float* latents, int B, int n_frames, float* out, size_t out_size) {
    const auto t0 = Clock::now();
    DeviceGuard guard(ctx->device);
    const VaeConfig& c = ctx->vae.cfg;
    const size_t expected = (size_t)B * n_frames * c.hop_length * c.audio_channels * sizeof(float);
    if (out_size < expected) return set_error(ctx, ACE_GGML_ERR_INVALID_ARG, "output buffer too small");
    const size_t lat_bytes = (size_t)B * n_frames * c.decoder_input_channels * 4;
    if (!ctx->io.reserve(lat_bytes, expected)) return set_error(ctx, ACE_GGML_ERR, "VAE I/O buffer allocation failed");
    const auto t_init = Clock::now();
    cudaError_t e = cudaMemcpyAsync(ctx->io.d_in, latents, lat_bytes, cudaMemcpyHostToDevice, ctx->stream);
    if (e != cudaSuccess) return cuda_fail(ctx, "latent upload failed", e);
    cudaEventRecord(ctx->ev0, ctx->stream);
    std::string err;
    e = vae_decode_grouped(ctx, static_cast<const float*>(ctx->io.d_in), B, n_frames, static_cast<float*>(ctx->io.d_out), err);
    cudaEventRecord(ctx->ev1, ctx->stream);
    if (e != cudaSuccess) return err.empty() ? cuda_fail(ctx, "vae decode launch failed", e) : set_error(ctx, ACE_GGML_ERR, err);
    const auto t_done_launch = Clock::now();
    (void)t_done_launch;
    e = cudaMemcpyAsync(out, ctx->io.d_out, expected, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) return cuda_fail(ctx, "vae decode failed", e);
    const auto t_done = Clock::now();
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->last_device_ms = ms;
    if (env_on("ACE_GGML_VAE_PROFILE")) {
        const auto t_end = Clock::now();
        std::fprintf(stderr,
                     "ace_ggml_vae_decode profile: backend=%s frames=%d out_len=%d threads=%d init_ms=%lld build_ms=%lld graph_ms=%lld "
                     "compute_ms=%lld copy_ms=%lld total_ms=%lld device_ms=%.3f batch=%d\n",
                     "b200", n_frames, n_frames * c.hop_length, ctx->n_threads, ms_between(t0, t_init), 0LL, 0LL,
                     ms_between(t_init, t_done), ms_between(t_done, t_end), ms_between(t0, t_end), ms, B);
    }
    return ACE_GGML_OK;
}

// Device-resident Euler sampler (+ optional CFG as one 2B-sample batch). Leaves the final latents in ctx->d_x.
ace_ggml_status sample_latents_device(ace_ggml_context* ctx, const float* noise, const float* context, const float* enc,
                                      const float* uncond_enc, const int32_t* mask, const int32_t* enc_mask, int B, int T,
                                      int E, const float* schedule, int n_steps, float guidance) {
    const DitConfig& c = ctx->dit.cfg;
    const int A = c.audio_acoustic_hidden_dim, cd = c.in_channels - A, H = c.hidden_size;
    const bool cfg = uncond_enc != nullptr && E > 0;
    const int EB = cfg ? 2 * B : B;  // evaluations per step
    std::string err;
    if (!dit_workspace_reserve(ctx->dit, ctx->dws, EB, T, E, err)) return set_error(ctx, ACE_GGML_ERR, err);
    DitWorkspace& ws = ctx->dws;
    const size_t n_lat = (size_t)B * T * A;
    if (ctx->sampler_cap < n_lat) {
        if (ctx->d_x) cudaFree(ctx->d_x);
        if (ctx->d_v) cudaFree(ctx->d_v);
        ctx->d_x = ctx->d_v = nullptr;
        ctx->sampler_cap = 0;
        if (cudaMalloc(&ctx->d_x, n_lat * 4) != cudaSuccess || cudaMalloc(&ctx->d_v, 2 * n_lat * 4) != cudaSuccess)
            return cuda_fail(ctx, "sampler cudaMalloc failed", cudaGetLastError());
        ctx->sampler_cap = n_lat;
    }
    cudaStream_t st = ctx->stream;
    cudaError_t e = cudaSuccess;
    auto chk = [&](cudaError_t r) { if (e == cudaSuccess) e = r; };
    chk(cudaMemcpyAsync(ctx->d_x, noise, n_lat * 4, cudaMemcpyHostToDevice, st));
    // step-invariant inputs: context, encoder states (cond | uncond), masks
    for (int rep = 0; rep < (cfg ? 2 : 1); ++rep) {
        float* dctx = ws.d_context + (size_t)rep * B * T * cd;
        if (context) chk(cudaMemcpyAsync(dctx, context, (size_t)B * T * cd * 4, cudaMemcpyHostToDevice, st));
        else chk(cudaMemsetAsync(dctx, 0, (size_t)B * T * cd * 4, st));
        if (E > 0) {
            const float* src = rep == 0 ? enc : uncond_enc;
            float* denc = ws.d_enc + (size_t)rep * B * E * H;
            if (src) chk(cudaMemcpyAsync(denc, src, (size_t)B * E * H * 4, cudaMemcpyHostToDevice, st));
            else chk(cudaMemsetAsync(denc, 0, (size_t)B * E * H * 4, st));
            if (enc_mask) chk(cudaMemcpyAsync(ws.d_enc_mask + (size_t)rep * B * E, enc_mask, (size_t)B * E * 4, cudaMemcpyHostToDevice, st));
        }
        if (mask) chk(cudaMemcpyAsync(ws.d_mask + (size_t)rep * B * T, mask, (size_t)B * T * 4, cudaMemcpyHostToDevice, st));
    }
    if (e != cudaSuccess) return cuda_fail(ctx, "sampler input upload failed", e);

    DitForwardOpts o;
    o.max_layers = env_max_layers();
    o.has_mask = mask != nullptr;
    o.has_enc_mask = enc_mask != nullptr && E > 0;
    ws.cross_kv_valid = false;
    std::vector<float> tt((size_t)2 * EB);
    // time embeddings of the whole schedule in one pass (every sample of the batch shares the step's t; t_r == t)
    const bool sched_ok = !env_on("ACE_B200_NO_SCHED_TEMB") && dit_time_embed_schedule(ctx->dit, ws, schedule, n_steps, st) == cudaSuccess;
    if (!sched_ok) cudaGetLastError();
    cudaEventRecord(ctx->ev0, st);
    for (int i = 0; i < n_steps; ++i) {
        const float t = schedule[i];
        for (int b = 0; b < EB; ++b) {
            tt[(size_t)b] = t;
            tt[(size_t)EB + b] = 0.f;  // timestep_r == timestep in every caller (acestep_ggml.cpp:2062-2071)
        }
        chk(cudaMemcpyAsync(ws.d_t, tt.data(), tt.size() * 4, cudaMemcpyHostToDevice, st));
        // current latents are the hidden_states of both CFG branches
        chk(cudaMemcpyAsync(ws.d_hidden, ctx->d_x, n_lat * 4, cudaMemcpyDeviceToDevice, st));
        if (cfg) chk(cudaMemcpyAsync(ws.d_hidden + n_lat, ctx->d_x, n_lat * 4, cudaMemcpyDeviceToDevice, st));
        o.reuse_cross_kv = i > 0;  // condition_embedder output and cross K/V do not depend on the step
        o.sched_step = sched_ok ? i : -1;
        chk(dit_forward_device(ctx->dit, ws, EB, T, E, ctx->d_v, o, st));
        const float dt = (i == n_steps - 1) ? t : (t - schedule[i + 1]);
        launch_euler_step(ctx->d_x, ctx->d_v, cfg ? ctx->d_v + n_lat : nullptr, guidance, dt, (long)n_lat, st);
        if (e != cudaSuccess) break;
    }
    cudaEventRecord(ctx->ev1, st);
    ws.cross_kv_valid = false;
    if (e != cudaSuccess) return cuda_fail(ctx, "sampler launch failed", e);
    return ACE_GGML_OK;
}

ace_ggml_status create_context(const ace_ggml_init_params* params, int dev, ace_ggml_context** out_ctx) {
    if (!out_ctx) return ACE_GGML_ERR_INVALID_ARG;
    *out_ctx = nullptr;
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev <= 0) {
        std::fprintf(stderr, "ace_ggml_create: no CUDA device available (%s); this library has no CPU path\n",
                     cudaGetErrorString(cudaGetLastError()));
        return ACE_GGML_ERR;
    }
    if (dev < 0) cudaGetDevice(&dev);
    if (dev < 0 || dev >= n_dev || dev >= ACE_MAX_DEVICES) return ACE_GGML_ERR_INVALID_ARG;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess) return ACE_GGML_ERR;
    if (prop.major != 10) {
        std::fprintf(stderr, "ace_ggml_create: device %d is sm_%d%d; this library is built for sm_100a only\n", dev, prop.major, prop.minor);
        return ACE_GGML_ERR_UNSUPPORTED;
    }
    ace_ggml_context* ctx = new (std::nothrow) ace_ggml_context();
    if (!ctx) return ACE_GGML_ERR;
    ctx->device = dev;
    ctx->n_threads = (params && params->n_threads > 0) ? params->n_threads : 1;
    DeviceGuard guard(dev);
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess || cudaEventCreate(&ctx->ev0) != cudaSuccess ||
        cudaEventCreate(&ctx->ev1) != cudaSuccess) {
        delete ctx;
        return ACE_GGML_ERR;
    }
    *out_ctx = ctx;
    return ACE_GGML_OK;
}

void release_peer_buffer(ace_ggml_context* ctx) {
    if (ctx->peer_buf) {
        if (ctx->peer_kind == 1) cudaFree(ctx->peer_buf);
        else if (ctx->peer_kind == 2) cudaIpcCloseMemHandle(ctx->peer_buf);
    }
    ctx->peer_buf = nullptr;
    ctx->peer_bytes = 0;
    ctx->peer_kind = 0;
}

}  // namespace ace_abi

using namespace ace_abi;

extern "C" {

ace_ggml_status ace_ggml_create(const ace_ggml_init_params* params, ace_ggml_context** out_ctx) {
    // device of the context: ACE_B200_DEVICE if set, else the calling thread's current CUDA device (ace_b200_create_on_device
    // names it explicitly, which is what a process that drives several GPUs from several threads should use)
    int dev = -1;
    if (const char* v = std::getenv("ACE_B200_DEVICE")) dev = std::atoi(v);
    return create_context(params, dev, out_ctx);
}

void ace_ggml_destroy(ace_ggml_context* ctx) {
    if (!ctx) return;
    {
        DeviceGuard guard(ctx->device);
        cudaStreamSynchronize(ctx->stream);
        ctx->dit.release();
        ctx->dws.release();
        ctx->vae.release();
        ctx->vws.release();
        ctx->vews.release();
        ctx->text.release();
        ctx->ews.release();
        ctx->io.release();
        release_peer_buffer(ctx);
        if (ctx->d_x) cudaFree(ctx->d_x);
        if (ctx->d_v) cudaFree(ctx->d_v);
        for (void* p : ctx->dbg_allocs) cudaFree(p);
        if (ctx->ev0) cudaEventDestroy(ctx->ev0);
        if (ctx->ev1) cudaEventDestroy(ctx->ev1);
        if (ctx->stream) {
            ace::decode_scratch_release(ctx->stream);
            cudaStreamDestroy(ctx->stream);
        }
    }
    delete ctx;
}

const char* ace_ggml_last_error(const ace_ggml_context* ctx) {
    if (!ctx) return "ace_ggml_last_error: null context";
    return ctx->last_error.c_str();
}

ace_ggml_status ace_ggml_load_dit(ace_ggml_context* ctx, const char* model_dir) {
    if (!ctx || !model_dir) return ACE_GGML_ERR_INVALID_ARG;
    DeviceGuard guard(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    ctx->dit_loaded = false;
    ctx->dws.release();
    std::string err;
    const int st = dit_load(model_dir, ctx->dit, err);
    if (st != 0) return set_error(ctx, map_load_status(st), err);
    ctx->dit_loaded = true;
    return ACE_GGML_OK;
}

ace_ggml_status ace_ggml_load_vae(ace_ggml_context* ctx, const char* model_dir) {
    if (!ctx || !model_dir) return ACE_GGML_ERR_INVALID_ARG;
    DeviceGuard guard(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    ctx->vae_loaded = false;
    ctx->vws.release();
    ctx->vews.release();
    std::string err;
    const int st = vae_load(model_dir, ctx->vae, err);
    if (st != 0) return set_error(ctx, map_load_status(st), err);
    ctx->vae_loaded = true;
    return ACE_GGML_OK;
}

ace_ggml_status ace_ggml_vae_get_info(ace_ggml_context* ctx, int32_t* latent_channels, int32_t* audio_channels, int32_t* hop_length) {
    if (!ctx) return ACE_GGML_ERR_INVALID_ARG;
    if (!ctx->vae_loaded) return set_error(ctx, ACE_GGML_ERR, "vae not loaded");
    if (latent_channels) *latent_channels = ctx->vae.cfg.decoder_input_channels;
    if (audio_channels) *audio_channels = ctx->vae.cfg.audio_channels;
    if (hop_length) *hop_length = ctx->vae.cfg.hop_length;
    return ACE_GGML_OK;
}

ace_ggml_status ace_ggml_vae_decode(ace_ggml_context* ctx, const float* latents, int32_t n_frames, float* out, size_t out_size) {
    if (!ctx || !latents || !out || n_frames <= 0) return ACE_GGML_ERR_INVALID_ARG;
    if (!ctx->vae_loaded) return set_error(ctx, ACE_GGML_ERR, "vae not loaded");
    return vae_decode_impl(ctx, latents, 1, n_frames, out, out_size);
}

ace_ggml_status ace_ggml_dit_forward(ace_ggml_context* ctx, const float* hidden_states, const float* context_latents,
                                     const float* encoder_hidden_states, const int32_t* attention_mask,
                                     const int32_t* encoder_attention_mask, int32_t seq_len, int32_t enc_len, float timestep,
                                     float timestep_r, float* out, size_t out_size) {
    if (!ctx || !out || seq_len <= 0) return ACE_GGML_ERR_INVALID_ARG;
    if (!ctx->dit_loaded) return set_error(ctx, ACE_GGML_ERR, "dit not loaded");
    if (enc_len < 0) enc_len = 0;
    return dit_forward_impl(ctx, hidden_states, context_latents, encoder_hidden_states, attention_mask, encoder_attention_mask, 1,
                            seq_len, enc_len, &timestep, &timestep_r, out, out_size, "ace_ggml_dit_forward");
}

}  // extern "C"
