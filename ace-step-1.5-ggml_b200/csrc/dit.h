// DiT velocity network: weights in HBM (packed tiles, see wpack.h) and the batched forward pass.
// reference: acestep_dit_config.h:10-32 (Config), acestep_dit_model.h:15-91 (weight handles),
//            acestep_dit_model.cpp:753-1088 (load_model_from_dir), 1316-1560 (forward_dit)
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include "encoders.h"
#include "wpack.h"

namespace ace {

struct DitConfig {
    int hidden_size = 0, intermediate_size = 0, num_hidden_layers = 0;
    int num_attention_heads = 0, num_key_value_heads = 0, head_dim = 0;
    int max_position_embeddings = 0;
    float rms_norm_eps = 1e-6f, rope_theta = 1000000.0f;
    int patch_size = 0, in_channels = 0, audio_acoustic_hidden_dim = 0;
    bool use_sliding_window = false;
    int sliding_window = 0;
    std::vector<std::string> layer_types;
    // condition encoders (optional keys, acestep_dit_config.cpp:74-78)
    int text_hidden_dim = 0, num_lyric_encoder_hidden_layers = 0, timbre_hidden_dim = 0, num_timbre_encoder_hidden_layers = 0;
    int timbre_fix_frame = 0;
};

struct DitLayer {
    float* self_norm = nullptr;   // [H]
    float* cross_norm = nullptr;  // [H]
    float* mlp_norm = nullptr;    // [H]
    PackedWeight self_qkv;        // rows: q (nq*hd) | k (nkv*hd) | v (nkv*hd)
    PackedWeight self_o;
    float* self_qn = nullptr;     // [hd]
    float* self_kn = nullptr;
    PackedWeight cross_q, cross_kv, cross_o;
    float* cross_qn = nullptr;
    float* cross_kn = nullptr;
    PackedWeight mlp_gate_up;     // rows interleaved (gate_j, up_j)
    PackedWeight mlp_down;
    bool sliding = false;
};

struct DitTimeEmbed {
    PackedWeight w1, w2, w_proj;
    float *b1 = nullptr, *b2 = nullptr, *b_proj = nullptr;
};

struct DitModel {
    DitConfig cfg;
    std::vector<DitLayer> layers;
    PackedWeight proj_in, proj_out, condition;
    float *proj_in_b = nullptr, *proj_out_b = nullptr /* expanded to [patch*audio_dim] */, *condition_b = nullptr;
    float* norm_out = nullptr;
    float* out_table = nullptr;   // [2, H]
    float* layer_tables = nullptr;  // [layers, 6, H]
    DitTimeEmbed te, te_r;
    // condition side (`encoder.*` tensors; all optional — acestep_dit_model.cpp:885-994): text projector, lyric / timbre encoders
    PackedWeight text_projector, lyric_embed, timbre_embed;
    bool has_text_projector = false, has_lyric_embed = false, has_timbre_embed = false;
    float *lyric_embed_b = nullptr, *timbre_embed_b = nullptr;
    EncStack lyric, timbre;
    std::vector<void*> allocs;    // every device allocation owned by the model
    size_t weight_bytes = 0;
    std::string weight_desc;      // e.g. "gguf:Q8_0" / "safetensors:F32->F16"

    void release();
};

// status: 0 ok, 3 IO (missing files / tensors), 4 unsupported, 1 other
int dit_load(const std::string& model_dir, DitModel& m, std::string& err);

// Per-context scratch, grown on demand.
struct DitWorkspace {
    int capB = 0, capT = 0, capE = 0;
    // inputs / outputs staged on the device
    float *d_hidden = nullptr, *d_context = nullptr, *d_enc = nullptr, *d_out = nullptr;
    int32_t *d_mask = nullptr, *d_enc_mask = nullptr, *d_pmask = nullptr;
    float* d_t = nullptr;  // [2, B]: t then (t - t_r)
    // activations
    __half *x0 = nullptr, *xn = nullptr, *qkv = nullptr, *att = nullptr, *act = nullptr;
    __half *enc16 = nullptr, *encp = nullptr, *cross_kv = nullptr /* [layers][B*E, 2*nkv*hd] */;
    float* x = nullptr;
    __half *tfreq = nullptr, *th1 = nullptr, *th2 = nullptr;
    float *temb = nullptr, *temb_r = nullptr, *tproj = nullptr, *mod = nullptr, *out_mod = nullptr;
    // schedule-wide time embeddings (device-resident sampler): [n_steps, H], [n_steps, H], [n_steps, 6H]
    float *sched_temb = nullptr, *sched_temb_r = nullptr, *sched_tproj = nullptr;
    __half *sched_freq = nullptr, *sched_h1 = nullptr, *sched_h2 = nullptr;
    float* sched_t = nullptr;
    int sched_cap = 0;
    float2* rope = nullptr;
    int rope_len = 0;
    bool cross_kv_valid = false;  // set by the device-resident sampler to reuse step-invariant K/V
    std::vector<void*> allocs;
    void release();
};

struct DitForwardOpts {
    int max_layers = 0;           // 0 = all (ACE_GGML_DIT_MAX_LAYERS, acestep_dit_model.cpp:1457-1464)
    bool reuse_cross_kv = false;  // skip condition_embedder + cross K/V projection (cached from a previous call)
    bool has_mask = false, has_enc_mask = false;
    // time embeddings precomputed for the whole schedule (dit_time_embed_schedule): row `step` of ws.sched_* is this call's, shared
    // by every sample of the batch; the six M = 1 GEMVs are then skipped
    int sched_step = -1;
};

// grows `ws` to hold (B, T, E); returns false + err on allocation failure
bool dit_workspace_reserve(const DitModel& m, DitWorkspace& ws, int B, int T, int E, std::string& err);

// Time-embedding MLPs of a whole schedule in one pass (every caller of the sampler has timestep_r == timestep, so the second embedder
// sees 0 at every step): the six weight-streaming M = 1 GEMVs per step become six launches with M = n_steps rows before the loop.
// t_host: [n_steps] timesteps.  reference: acestep_dit_model.cpp:1286-1308, 1420-1424
cudaError_t dit_time_embed_schedule(const DitModel& m, DitWorkspace& ws, const float* t_host, int n_steps, cudaStream_t st);

// Batched forward on device-resident inputs already placed in ws.d_hidden [B,T,A], ws.d_context [B,T,ctx],
// ws.d_enc [B,E,H], ws.d_mask [B,T], ws.d_enc_mask [B,E], ws.d_t [2,B]. Writes velocity to `out` [B,T,A] (device).
cudaError_t dit_forward_device(const DitModel& m, DitWorkspace& ws, int B, int T, int E, float* out,
                               const DitForwardOpts& opts, cudaStream_t st);

}  // namespace ace
