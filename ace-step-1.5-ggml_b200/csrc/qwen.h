// Qwen3 text encoder (Qwen3-Embedding-0.6B shape: causal GQA transformer with per-head q/k RMSNorm and NEOX RoPE) in HBM.
// reference: qwen_config.h:10-22 / qwen_config.cpp:19-70 (keys), qwen_model.cpp:47-71 (file choice), 336-478 (tensor names),
//            528-695 (forward_text_encoder_layers / _embeddings)
#pragma once
#include <string>
#include <vector>

#include "encoders.h"

namespace ace {

struct QwenConfig {
    int vocab_size = 0, hidden_size = 0, num_hidden_layers = 0, num_attention_heads = 0, num_key_value_heads = 0;
    int intermediate_size = 0, head_dim = 0, max_position_embeddings = 0;
    float rms_norm_eps = 0.f, rope_theta = 1000000.0f;
};

struct QwenModel {
    QwenConfig cfg;
    EncStack stack;
    float* table = nullptr;  // [vocab, hidden] fp32: the file's rows dequantised exactly as ggml_get_rows + cast would (qwen_model.cpp:553-554)
    std::vector<void*> allocs;
    size_t weight_bytes = 0;
    bool loaded = false;
    std::string weight_desc;
    void release();
};

// status: 0 ok, 3 IO (missing files / tensors), 4 unsupported, 1 other
int qwen_load(const std::string& model_dir, QwenModel& m, std::string& err);

}  // namespace ace
