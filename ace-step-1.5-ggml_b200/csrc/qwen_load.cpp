// Qwen3 text-encoder ingest: config.json + (model.gguf | model.safetensors) -> packed tiles + fp32 embedding table in HBM.
// reference: qwen_config.cpp:19-70, qwen_model.cpp:27-45 (on-the-fly quantisation switch), 47-71 (GGUF resolution), 336-478
#include "qwen.h"

#include <algorithm>
#include <cstdlib>

namespace ace {

void QwenModel::release() {
    for (void* p : allocs) cudaFree(p);
    allocs.clear();
    stack = EncStack();
    table = nullptr;
    weight_bytes = 0;
    loaded = false;
}

namespace {

bool parse_qwen_config(const std::string& path, QwenConfig& c, std::string& err) {
    JsonValue root;
    if (!json_parse_file(path, root, err)) return false;
    if (root.kind != JsonValue::Object) {
        err = "config is not object";
        return false;
    }
    auto need = [&](const char* k, int& dst) {
        int64_t v;
        if (!root.get_int(k, v)) {
            if (err.empty()) err = std::string("config.json: missing key ") + k;
            return false;
        }
        dst = (int)v;
        return true;
    };
    if (!need("vocab_size", c.vocab_size) || !need("hidden_size", c.hidden_size) || !need("num_hidden_layers", c.num_hidden_layers) ||
        !need("num_attention_heads", c.num_attention_heads) || !need("num_key_value_heads", c.num_key_value_heads) ||
        !need("intermediate_size", c.intermediate_size) || !need("head_dim", c.head_dim) ||
        !need("max_position_embeddings", c.max_position_embeddings))
        return false;
    double d;
    if (!root.get_num("rms_norm_eps", d)) {
        err = "config.json: missing key rms_norm_eps";
        return false;
    }
    c.rms_norm_eps = (float)d;
    if (root.get_num("rope_theta", d)) c.rope_theta = (float)d;
    return true;
}

std::string resolve_qwen_gguf(const std::string& dir) {
    for (const char* key : {"ACE_GGML_QWEN_GGUF", "ACE_GGML_TEXT_ENCODER_GGUF", "ACE_GGML_LM_GGUF"}) {
        const char* v = std::getenv(key);
        if (v && v[0] && file_exists(v)) return v;
    }
    if (dir.size() > 5 && dir.substr(dir.size() - 5) == ".gguf" && file_exists(dir)) return dir;
    if (file_exists(dir + "/model.gguf")) return dir + "/model.gguf";
    return "";
}

// W_Q8_0 / W_Q4_K / W_Q6_K, or -1 (nothing asked for)
int qwen_online_qtype(std::string& unsupported) {
    for (const char* key : {"ACE_GGML_QWEN_WEIGHT_QTYPE", "ACE_GGML_WEIGHT_QTYPE"}) {
        const char* v = std::getenv(key);
        if (!v || !v[0]) continue;
        std::string s(v);
        std::transform(s.begin(), s.end(), s.begin(), [](unsigned char ch) { return (char)std::toupper(ch); });
        if (s == "Q8" || s == "Q8_0") return W_Q8_0;
        if (s == "Q6" || s == "Q6_K") return W_Q6_K;
        if (s == "Q4" || s == "Q4_K" || s == "Q4_K_M") return W_Q4_K;
        (void)unsupported;
    }
    return -1;
}

}  // namespace

int qwen_load(const std::string& model_dir, QwenModel& m, std::string& err) {
    m.release();
    m.cfg = QwenConfig();
    std::string root = model_dir;
    if (root.size() > 5 && root.substr(root.size() - 5) == ".gguf") {
        const size_t slash = root.find_last_of('/');
        root = slash == std::string::npos ? "." : root.substr(0, slash);
    }
    if (!parse_qwen_config(root + "/config.json", m.cfg, err)) return 3;
    const QwenConfig& c = m.cfg;
    if (c.head_dim != 128) {
        err = "unsupported head_dim (the attention kernel is specialised on 128)";
        return 4;
    }
    if (c.hidden_size % 64 != 0 || c.intermediate_size % 64 != 0 || c.num_key_value_heads <= 0 ||
        c.num_attention_heads % c.num_key_value_heads != 0 || c.vocab_size <= 0 || c.num_hidden_layers < 0) {
        err = "unsupported text-encoder dimensions";
        return 4;
    }
    std::unique_ptr<TensorFile> tf;
    const std::string gguf = resolve_qwen_gguf(model_dir);
    if (!gguf.empty()) tf = TensorFile::open_gguf(gguf, err);
    else tf = TensorFile::open_safetensors(root + "/model.safetensors", err);
    if (!tf) return 3;

    WeightLoader L{tf.get(), &m.allocs, &m.weight_bytes, {}, 0, -1};
    std::string unsup;
    L.online_q = qwen_online_qtype(unsup);
    EncStack& st = m.stack;
    st.H = c.hidden_size;
    st.I = c.intermediate_size;
    st.n_heads = c.num_attention_heads;
    st.n_kv = c.num_key_value_heads;
    st.hd = c.head_dim;
    st.eps = c.rms_norm_eps;
    st.theta = c.rope_theta;
    st.sliding_window = 0;

    bool ok = true;
    {  // embedding table: every row decoded to fp32 exactly as ggml's get_rows dequantiser does
        std::vector<float> tab;
        std::vector<int64_t> sh;
        std::string e;
        if (!tf->read_f32("embed_tokens.weight", tab, &sh, e)) ok = L.fail(3, e);
        else if (sh.size() != 2 || sh[1] != c.hidden_size || sh[0] < 1) ok = L.fail(3, "invalid tensor shape for embed_tokens.weight");
        else {
            if (sh[0] < c.vocab_size) m.cfg.vocab_size = (int)sh[0];
            const int oq = tf->is_gguf() ? -1 : L.online_q;
            if ((oq == W_Q8_0 && c.hidden_size % 32 == 0) || ((oq == W_Q4_K || oq == W_Q6_K) && c.hidden_size % 256 == 0)) {
                // the reference quantises the table too when the on-line switch is set (qwen_model.cpp:124-146): round-trip it
                const int gt = oq == W_Q8_0 ? GGML_Q8_0 : (oq == W_Q4_K ? GGML_Q4_K : GGML_Q6_K);
                std::vector<uint8_t> q(ggml_row_bytes(gt, c.hidden_size));
                for (int64_t r = 0; r < sh[0]; ++r) {
                    float* row = tab.data() + r * c.hidden_size;
                    if (gt == GGML_Q8_0) quantize_row_q8_0(row, q.data(), c.hidden_size);
                    else if (gt == GGML_Q4_K) quantize_row_q4_k(row, q.data(), c.hidden_size);
                    else quantize_row_q6_k(row, q.data(), c.hidden_size);
                    dequant_row(gt, q.data(), row, c.hidden_size);
                }
            }
            m.table = L.upload_f32(tab);
            ok = m.table != nullptr;
        }
    }
    ok = ok && L.vec("norm.weight", c.hidden_size, st.final_norm);
    ok = ok && enc_stack_load(L, "layers.", c.num_hidden_layers, st);
    if (!ok) {
        err = L.err;
        m.release();
        return L.status ? L.status : 1;
    }
    if (cudaDeviceSynchronize() != cudaSuccess) {  // predecode kernels ran on the default stream
        err = std::string("weight predecode failed: ") + cudaGetErrorString(cudaGetLastError());
        m.release();
        return 1;
    }
    const int wt = st.layers.empty() ? W_F16 : st.layers[0].qkv.wtype;
    m.weight_desc = std::string(tf->is_gguf() ? "gguf:" : "safetensors:") + (wt == W_F16 ? "F16" : wt == W_Q8_0 ? "Q8_0" : wt == W_Q4_K ? "Q4_K" : "Q6_K");
    m.loaded = true;
    return 0;
}

}  // namespace ace
