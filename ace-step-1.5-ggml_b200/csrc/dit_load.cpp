// DiT weight ingest: config.json + (model.gguf | model.safetensors) -> packed tiles in HBM.
// reference: acestep_dit_config.cpp:19-93 (keys), acestep_dit_model.cpp:47-70 (file choice),
//            156-192 (on-the-fly quantisation), 334-490 / 640-715 (conv -> linear relayout), 753-1088 (tensor names)
#include "dit.h"
#include "gemm.h"
#include "tensor_file.h"
#include "loader.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace ace {

namespace {

using Loader = WeightLoader;

bool parse_config(const std::string& path, DitConfig& c, std::string& err) {
    JsonValue root;
    if (!json_parse_file(path, root, err)) return false;
    if (root.kind != JsonValue::Object) {
        err = "config is not object";
        return false;
    }
    auto need_int = [&](const char* k, int& dst) {
        int64_t v;
        if (!root.get_int(k, v)) {
            if (err.empty()) err = std::string("config.json: missing key ") + k;
            return false;
        }
        dst = (int)v;
        return true;
    };
    double d;
    int64_t iv;
    if (!need_int("hidden_size", c.hidden_size) || !need_int("intermediate_size", c.intermediate_size) ||
        !need_int("num_hidden_layers", c.num_hidden_layers) || !need_int("num_attention_heads", c.num_attention_heads) ||
        !need_int("num_key_value_heads", c.num_key_value_heads) || !need_int("head_dim", c.head_dim) ||
        !need_int("max_position_embeddings", c.max_position_embeddings))
        return false;
    if (!root.get_num("rms_norm_eps", d)) {
        err = "config.json: missing key rms_norm_eps";
        return false;
    }
    c.rms_norm_eps = (float)d;
    if (!need_int("patch_size", c.patch_size) || !need_int("in_channels", c.in_channels) ||
        !need_int("audio_acoustic_hidden_dim", c.audio_acoustic_hidden_dim))
        return false;
    if (root.get_int("text_hidden_dim", iv)) c.text_hidden_dim = (int)iv;
    if (root.get_int("num_lyric_encoder_hidden_layers", iv)) c.num_lyric_encoder_hidden_layers = (int)iv;
    if (root.get_int("timbre_hidden_dim", iv)) c.timbre_hidden_dim = (int)iv;
    if (root.get_int("num_timbre_encoder_hidden_layers", iv)) c.num_timbre_encoder_hidden_layers = (int)iv;
    if (root.get_int("timbre_fix_frame", iv)) c.timbre_fix_frame = (int)iv;
    (void)root.get_bool("use_sliding_window", c.use_sliding_window);
    if (root.get_int("sliding_window", iv)) c.sliding_window = (int)iv;
    if (root.get_num("rope_theta", d)) c.rope_theta = (float)d;
    const JsonValue* lt = root.get("layer_types");
    if (!lt || lt->kind != JsonValue::Array) {
        err = "missing layer_types";
        return false;
    }
    for (const auto& v : lt->arr) c.layer_types.push_back(v.str);
    return true;
}

std::string resolve_gguf(const std::string& dir) {
    for (const char* key : {"ACE_GGML_DIT_GGUF", "ACE_GGML_DIT_GGUF_PATH"}) {
        const char* v = std::getenv(key);
        if (v && v[0] && file_exists(v)) return v;
    }
    if (dir.size() > 5 && dir.substr(dir.size() - 5) == ".gguf" && file_exists(dir)) return dir;
    const std::string cand = dir + "/model.gguf";
    if (file_exists(cand)) return cand;
    return "";
}

int parse_online_qtype(std::string& unsupported) {
    for (const char* key : {"ACE_GGML_DIT_WEIGHT_QTYPE", "ACE_GGML_WEIGHT_QTYPE"}) {
        const char* v = std::getenv(key);
        if (!v || !v[0]) continue;
        std::string s(v);
        std::transform(s.begin(), s.end(), s.begin(), [](unsigned char ch) { return (char)std::toupper(ch); });
        if (s == "Q8" || s == "Q8_0") return W_Q8_0;
        if (s == "Q6" || s == "Q6_K") return W_Q6_K;
        if (s == "Q4" || s == "Q4_K" || s == "Q4_K_M") return W_Q4_K;
        (void)unsupported;
        return -1;  // unknown strings are ignored by the reference too
    }
    return -1;
}

}  // namespace

void DitModel::release() {
    for (void* p : allocs) cudaFree(p);
    allocs.clear();
    layers.clear();
    lyric = EncStack();
    timbre = EncStack();
    has_text_projector = has_lyric_embed = has_timbre_embed = false;
    lyric_embed_b = timbre_embed_b = nullptr;
    weight_bytes = 0;
}

int dit_load(const std::string& model_dir, DitModel& m, std::string& err) {
    m.release();
    m.cfg = DitConfig();
    std::string cfg_dir = model_dir;
    if (cfg_dir.size() > 5 && cfg_dir.substr(cfg_dir.size() - 5) == ".gguf") {
        const size_t slash = cfg_dir.find_last_of('/');
        cfg_dir = slash == std::string::npos ? "." : cfg_dir.substr(0, slash);
    }
    if (!parse_config(cfg_dir + "/config.json", m.cfg, err)) return 3;
    const DitConfig& c = m.cfg;
    if (c.head_dim != 128) {
        err = "unsupported head_dim (the attention kernel is specialised on 128)";
        return 4;
    }
    if ((int)c.layer_types.size() < c.num_hidden_layers) {
        err = "layer_types shorter than num_hidden_layers";
        return 3;
    }
    if (c.hidden_size % 64 != 0 || c.intermediate_size % 64 != 0 || c.num_attention_heads % c.num_key_value_heads != 0) {
        err = "unsupported DiT dimensions";
        return 4;
    }

    std::unique_ptr<TensorFile> tf;
    const std::string gguf = resolve_gguf(model_dir);
    if (!gguf.empty()) tf = TensorFile::open_gguf(gguf, err);
    else tf = TensorFile::open_safetensors(model_dir + "/model.safetensors", err);
    if (!tf) return 3;

    Loader L{tf.get(), &m.allocs, &m.weight_bytes, {}, 0, -1};
    std::string unsup;
    L.online_q = parse_online_qtype(unsup);

    const int H = c.hidden_size, I = c.intermediate_size, hd = c.head_dim;
    const int nq = c.num_attention_heads * hd, nkv = c.num_key_value_heads * hd;
    const int P = c.patch_size, C = c.in_channels, A = c.audio_acoustic_hidden_dim;
    bool ok = true;
    std::string e;

    // proj_in: conv1d weight [H, C, P] -> linear [H, P*C] with column index k*C + c   (acestep_dit_model.cpp:366-375)
    {
        std::vector<float> w;
        std::vector<int64_t> sh;
        if (!tf->read_f32("decoder.proj_in.1.weight", w, &sh, e)) ok = L.fail(3, e);
        else if (sh.size() != 3 || sh[0] != H || sh[1] != C || sh[2] != P) ok = L.fail(3, "invalid tensor shape for decoder.proj_in.1.weight");
        else {
            std::vector<float> lin((size_t)H * P * C);
            for (int o = 0; o < H; ++o)
                for (int ci = 0; ci < C; ++ci)
                    for (int k = 0; k < P; ++k) lin[(size_t)o * P * C + (size_t)k * C + ci] = w[((size_t)o * C + ci) * P + k];
            std::vector<uint8_t> img;
            ok = ok && L.pack_f32(lin, H, P * C, img, m.proj_in) && L.upload(m.proj_in, img);
        }
        ok = ok && L.vec("decoder.proj_in.1.bias", H, m.proj_in_b);
    }
    // proj_out: conv-transpose weight [H, A, P] -> linear [P*A, H] with row index k*A + a   (acestep_dit_model.cpp:445-454)
    if (ok) {
        std::vector<float> w, b;
        std::vector<int64_t> sh;
        if (!tf->read_f32("decoder.proj_out.1.weight", w, &sh, e)) ok = L.fail(3, e);
        else if (sh.size() != 3 || sh[0] != H || sh[1] != A || sh[2] != P) ok = L.fail(3, "invalid tensor shape for decoder.proj_out.1.weight");
        else {
            std::vector<float> lin((size_t)P * A * H);
            for (int i = 0; i < H; ++i)
                for (int a = 0; a < A; ++a)
                    for (int k = 0; k < P; ++k) lin[((size_t)k * A + a) * H + i] = w[((size_t)i * A + a) * P + k];
            std::vector<uint8_t> img;
            ok = ok && L.pack_f32(lin, P * A, H, img, m.proj_out) && L.upload(m.proj_out, img);
        }
        if (ok && !tf->read_f32("decoder.proj_out.1.bias", b, nullptr, e)) ok = L.fail(3, e);
        if (ok && (int)b.size() != A) ok = L.fail(3, "unexpected size for decoder.proj_out.1.bias");
        if (ok) {
            std::vector<float> be((size_t)P * A);
            for (int k = 0; k < P; ++k)
                for (int a = 0; a < A; ++a) be[(size_t)k * A + a] = b[(size_t)a];
            m.proj_out_b = L.upload_f32(be);
            ok = m.proj_out_b != nullptr;
        }
    }
    ok = ok && L.linear({"decoder.condition_embedder.weight"}, false, H, m.condition);
    ok = ok && L.vec("decoder.condition_embedder.bias", H, m.condition_b);
    ok = ok && L.vec("decoder.norm_out.weight", H, m.norm_out);
    if (ok) {
        std::vector<float> t;
        if (!tf->read_f32("decoder.scale_shift_table", t, nullptr, e)) ok = L.fail(3, e);
        else if ((int64_t)t.size() != 2L * H) ok = L.fail(3, "unexpected size for decoder.scale_shift_table");
        else {
            L.online_roundtrip(t, H);
            m.out_table = L.upload_f32(t);
            ok = m.out_table != nullptr;
        }
    }
    auto load_te = [&](const std::string& p, DitTimeEmbed& te) {
        return L.linear({p + ".linear_1.weight"}, false, 256, te.w1) && L.vec(p + ".linear_1.bias", H, te.b1) &&
               L.linear({p + ".linear_2.weight"}, false, H, te.w2) && L.vec(p + ".linear_2.bias", H, te.b2) &&
               L.linear({p + ".time_proj.weight"}, false, H, te.w_proj) && L.vec(p + ".time_proj.bias", 6L * H, te.b_proj);
    };
    ok = ok && load_te("decoder.time_embed", m.te) && load_te("decoder.time_embed_r", m.te_r);

    std::vector<float> tables((size_t)c.num_hidden_layers * 6 * H);
    m.layers.resize((size_t)c.num_hidden_layers);
    for (int i = 0; ok && i < c.num_hidden_layers; ++i) {
        DitLayer& ly = m.layers[(size_t)i];
        const std::string p = "decoder.layers." + std::to_string(i) + ".";
        ly.sliding = c.layer_types[(size_t)i] == "sliding_attention";
        ok = ok && L.vec(p + "self_attn_norm.weight", H, ly.self_norm) && L.vec(p + "cross_attn_norm.weight", H, ly.cross_norm) &&
             L.vec(p + "mlp_norm.weight", H, ly.mlp_norm);
        ok = ok && L.linear({p + "self_attn.q_proj.weight", p + "self_attn.k_proj.weight", p + "self_attn.v_proj.weight"}, false, H, ly.self_qkv);
        ok = ok && L.linear({p + "self_attn.o_proj.weight"}, false, nq, ly.self_o);
        ok = ok && L.vec(p + "self_attn.q_norm.weight", hd, ly.self_qn) && L.vec(p + "self_attn.k_norm.weight", hd, ly.self_kn);
        ok = ok && L.linear({p + "cross_attn.q_proj.weight"}, false, H, ly.cross_q);
        ok = ok && L.linear({p + "cross_attn.k_proj.weight", p + "cross_attn.v_proj.weight"}, false, H, ly.cross_kv);
        ok = ok && L.linear({p + "cross_attn.o_proj.weight"}, false, nq, ly.cross_o);
        ok = ok && L.vec(p + "cross_attn.q_norm.weight", hd, ly.cross_qn) && L.vec(p + "cross_attn.k_norm.weight", hd, ly.cross_kn);
        ok = ok && L.linear({p + "mlp.gate_proj.weight", p + "mlp.up_proj.weight"}, true, H, ly.mlp_gate_up);
        ok = ok && L.linear({p + "mlp.down_proj.weight"}, false, I, ly.mlp_down);
        if (ok) {
            std::vector<float> t;
            if (!tf->read_f32(p + "scale_shift_table", t, nullptr, e)) ok = L.fail(3, e);
            else if ((int64_t)t.size() != 6L * H) ok = L.fail(3, "unexpected size for " + p + "scale_shift_table");
            else {
                L.online_roundtrip(t, H);
                std::memcpy(tables.data() + (size_t)i * 6 * H, t.data(), t.size() * 4);
            }
        }
        if (ok && (ly.self_qkv.N != nq + 2 * nkv || ly.self_o.N != H || ly.cross_q.N != nq || ly.cross_kv.N != 2 * nkv ||
                   ly.mlp_gate_up.N != 2 * I || ly.mlp_down.N != H))
            ok = L.fail(3, "unexpected projection shape in layer " + std::to_string(i));
    }
    if (ok) {
        m.layer_tables = L.upload_f32(tables);
        ok = m.layer_tables != nullptr;
    }
    // ---- condition side: `encoder.*` (acestep_dit_model.cpp:885-994).  Projections are optional (loaded when present); the
    // layer stacks are required as soon as config.json announces them, exactly like the reference loader.
    auto init_stack = [&](EncStack& st) {
        st.H = H;
        st.I = I;
        st.n_heads = c.num_attention_heads;
        st.n_kv = c.num_key_value_heads;
        st.hd = hd;
        st.eps = c.rms_norm_eps;
        st.theta = c.rope_theta;
        st.sliding_window = c.sliding_window;
    };
    auto opt_linear = [&](const std::string& name, PackedWeight& w, bool& has) {
        const TensorInfo* t = tf->find(name);
        has = false;
        if (!t) return true;
        if (t->shape.size() != 2 || t->shape[1] % 8 != 0) return L.fail(3, "invalid tensor shape for " + name);
        has = L.linear({name}, false, (int)t->shape[1], w);
        return has;
    };
    auto opt_vec = [&](const std::string& name, int64_t n, float*& dst) { return !tf->find(name) || L.vec(name, n, dst); };
    ok = ok && opt_linear("encoder.text_projector.weight", m.text_projector, m.has_text_projector);
    ok = ok && opt_linear("encoder.lyric_encoder.embed_tokens.weight", m.lyric_embed, m.has_lyric_embed);
    ok = ok && opt_vec("encoder.lyric_encoder.embed_tokens.bias", H, m.lyric_embed_b);
    init_stack(m.lyric);
    init_stack(m.timbre);
    ok = ok && opt_vec("encoder.lyric_encoder.norm.weight", H, m.lyric.final_norm);
    if (ok && c.num_lyric_encoder_hidden_layers > 0) {
        ok = enc_stack_load(L, "encoder.lyric_encoder.layers.", c.num_lyric_encoder_hidden_layers, m.lyric);
        for (int i = 0; ok && i < c.num_lyric_encoder_hidden_layers; ++i)  // layer_types of the decoder index the encoder layers too (:933-935)
            m.lyric.layers[(size_t)i].sliding = i < (int)c.layer_types.size() && c.layer_types[(size_t)i] == "sliding_attention";
    }
    ok = ok && opt_linear("encoder.timbre_encoder.embed_tokens.weight", m.timbre_embed, m.has_timbre_embed);
    ok = ok && opt_vec("encoder.timbre_encoder.embed_tokens.bias", H, m.timbre_embed_b);
    ok = ok && opt_vec("encoder.timbre_encoder.norm.weight", H, m.timbre.final_norm);
    if (ok && c.num_timbre_encoder_hidden_layers > 0) {
        ok = enc_stack_load(L, "encoder.timbre_encoder.layers.", c.num_timbre_encoder_hidden_layers, m.timbre);
        for (int i = 0; ok && i < c.num_timbre_encoder_hidden_layers; ++i)
            m.timbre.layers[(size_t)i].sliding = i < (int)c.layer_types.size() && c.layer_types[(size_t)i] == "sliding_attention";
    }
    if (!ok) {
        err = L.err;
        m.release();
        return L.status ? L.status : 1;
    }
    m.weight_desc = std::string(tf->is_gguf() ? "gguf:" : "safetensors:") +
                    (m.layers.empty() ? "?" : (m.layers[0].self_qkv.wtype == W_F16 ? "F16" : m.layers[0].self_qkv.wtype == W_Q8_0 ? "Q8_0"
                                               : m.layers[0].self_qkv.wtype == W_Q4_K ? "Q4_K" : "Q6_K"));
    if (!m.layers.empty() && m.layers[0].self_qkv.dev_f16) m.weight_desc += "+f16twin";
    if (cudaDeviceSynchronize() != cudaSuccess) {  // predecode kernels ran on the default stream
        err = std::string("weight predecode failed: ") + cudaGetErrorString(cudaGetLastError());
        m.release();
        return 1;
    }
    return 0;
}

}  // namespace ace
