// Plain pre-norm transformer stack (no AdaLN, no cross-attention) on the same kernels as the DiT: the lyric and timbre
// condition encoders and the Qwen3 text encoder are all this block with different weights and mask modes.
//   layer: x += o( attn( headnorm+rope( qkv( rmsnorm(x) ) ) ) );  x += down( silu(gate) * up )( rmsnorm(x) )
// reference: acestep_dit_model.cpp:1562-1739 (forward_lyric_encoder / forward_timbre_encoder: bidirectional, RoPE, optional
//            sliding window per layer_types), qwen_model.cpp:528-666 (forward_text_encoder_layers: causal + key-padding mask)
#pragma once
#include <string>
#include <vector>

#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include "loader.h"
#include "wpack.h"

namespace ace {

struct EncLayer {
    float* input_norm = nullptr;  // [H]
    float* post_norm = nullptr;   // [H]
    PackedWeight qkv, o, gate_up, down;
    float* qn = nullptr;          // [head_dim]
    float* kn = nullptr;
    bool sliding = false;
};

struct EncStack {
    int H = 0, I = 0, n_heads = 0, n_kv = 0, hd = 128;
    float eps = 1e-6f, theta = 1000000.0f;
    int sliding_window = 0;
    std::vector<EncLayer> layers;
    float* final_norm = nullptr;  // [H] or null
};

// loads layers `prefix + "<i>." + {input_layernorm, self_attn.{q,k,v,o}_proj, self_attn.{q,k}_norm, post_attention_layernorm,
// mlp.{gate,up,down}_proj}` (the names both reference loaders use: acestep_dit_model.cpp:902-937, qwen_model.cpp:436-470)
bool enc_stack_load(WeightLoader& L, const std::string& prefix, int n_layers, EncStack& st);

struct EncWorkspace {
    int cap_rows = 0;
    float* x = nullptr;                                       // fp32 residual stream [rows, H]
    __half *xn = nullptr, *qkv = nullptr, *att = nullptr, *act = nullptr;
    float2* rope = nullptr;
    int rope_len = 0;
    float rope_theta = 0.f;
    int32_t* d_mask = nullptr;
    size_t cap_elems = 0;                                     // capacity is tracked in (rows x widest row) elements
    std::vector<void*> allocs;
    void release();
};
bool enc_workspace_reserve(const EncStack& st, EncWorkspace& ws, int rows, std::string& err);

// One sample of `rows` tokens whose fp32 embeddings are already in ws.x.  Runs the first n_layers layers (<0 or > size = all),
// then the final norm when `final_norm` is set (and the stack has one); writes fp32 [rows, H] to `out` (device).
// key_mask (device, [rows], may be null) hides padded keys; causal hides keys after the query.
cudaError_t enc_stack_forward(const EncStack& st, EncWorkspace& ws, int rows, int n_layers, bool causal, const int32_t* key_mask,
                              bool final_norm, float* out, cudaStream_t stream);

}  // namespace ace
