// Oobleck VAE decode on sm_100a: every conv / conv-transpose is one tap-GEMM (gemm.h) with bias, Snake and the
// residual add fused in its epilogue; activations are channels-last [L, C] (fp16 operands, fp32 residual stream).
// reference: acestep_vae_model.cpp:520-588 (weight-norm fold, F16 store), 682-742 (snake / conv / residual / block),
//            957-1002 (forward_decode);  ggml_conv_1d = F16 im2col + F16 matmul (ggml.c:4337-4426)
#include "vae.h"
#include "gemm.h"
#include "kernels.h"
#include "tensor_file.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>

namespace ace {

void VaeModel::release() {
    for (void* p : allocs) cudaFree(p);
    allocs.clear();
    blocks.clear();
    enc_blocks.clear();
    has_encoder = false;
    weight_bytes = 0;
}
void VaeEncWorkspace::release() {
    for (void* p : allocs) cudaFree(p);
    allocs.clear();
    *this = VaeEncWorkspace();
}
void VaeWorkspace::release() {
    for (void* p : allocs) cudaFree(p);
    allocs.clear();
    *this = VaeWorkspace();
}

namespace {

struct VLoader {
    const TensorFile* tf;
    VaeModel* m;
    std::string err;
    int status = 0;

    bool fail(int st, const std::string& e) {
        if (status == 0) { status = st; err = e; }
        return false;
    }
    void* dev_upload(const void* src, size_t bytes) {
        void* p = nullptr;
        if (cudaMalloc(&p, bytes ? bytes : 16) != cudaSuccess) {
            fail(1, std::string("cudaMalloc failed: ") + cudaGetErrorString(cudaGetLastError()));
            return nullptr;
        }
        m->allocs.push_back(p);
        m->weight_bytes += bytes;
        if (cudaMemcpy(p, src, bytes, cudaMemcpyHostToDevice) != cudaSuccess) {
            fail(1, std::string("cudaMemcpy failed: ") + cudaGetErrorString(cudaGetLastError()));
            return nullptr;
        }
        return p;
    }
    // w = v * g / sqrt(sum_row(v^2) + 1e-12) per d0 row   (acestep_vae_model.cpp:552-565)
    bool folded(const std::string& prefix, std::vector<float>& w, std::vector<int64_t>& shape) {
        std::vector<float> g, v;
        std::vector<int64_t> gs;
        std::string e;
        if (!tf->read_f32(prefix + ".weight_g", g, &gs, e) || !tf->read_f32(prefix + ".weight_v", v, &shape, e)) return fail(3, e);
        if (shape.size() != 3 || gs.size() != 3) return fail(3, "invalid weight-norm tensor shape for " + prefix);
        const int64_t d0 = shape[0], row = shape[1] * shape[2];
        if ((int64_t)g.size() != d0) return fail(3, "weight_g size mismatch for " + prefix);
        w.resize(v.size());
        for (int64_t i = 0; i < d0; ++i) {
            double ss = 0.0;
            for (int64_t j = 0; j < row; ++j) ss += (double)v[(size_t)(i * row + j)] * (double)v[(size_t)(i * row + j)];
            const float scale = g[(size_t)i] / std::sqrt((float)ss + 1e-12f);
            for (int64_t j = 0; j < row; ++j) w[(size_t)(i * row + j)] = v[(size_t)(i * row + j)] * scale;
        }
        return true;
    }
    bool bias_vec(const std::string& name, int n, int repeat, float*& out) {
        std::vector<float> b;
        std::string e;
        if (!tf->read_f32(name, b, nullptr, e)) return fail(3, e);
        if ((int)b.size() != n) return fail(3, "unexpected size for " + name);
        std::vector<float> t((size_t)n * repeat);
        for (int r = 0; r < repeat; ++r) std::memcpy(t.data() + (size_t)r * n, b.data(), (size_t)n * 4);
        out = static_cast<float*>(dev_upload(t.data(), t.size() * 4));
        return out != nullptr;
    }
    bool snake(const std::string& prefix, int C, VaeSnake& s) {
        std::vector<float> a, b;
        std::string e;
        if (!tf->read_f32(prefix + ".alpha", a, nullptr, e) || !tf->read_f32(prefix + ".beta", b, nullptr, e)) return fail(3, e);
        if ((int)a.size() != C || (int)b.size() != C) return fail(3, "unexpected snake size for " + prefix);
        for (auto& x : a) x = std::exp(x);
        for (auto& x : b) x = 1.0f / std::exp(x);  // no +1e-9: the ggml graph divides by exp(beta) directly (:682-692)
        s.alpha = static_cast<float*>(dev_upload(a.data(), a.size() * 4));
        s.inv_beta = static_cast<float*>(dev_upload(b.data(), b.size() * 4));
        return s.alpha && s.inv_beta;
    }
    // plain conv: weight_v [Cout, Cin, K] -> W'[co][tap][ci]
    bool conv(const std::string& prefix, int dilation, bool with_bias, VaeConv& c) {
        std::vector<float> w;
        std::vector<int64_t> sh;
        if (!folded(prefix, w, sh)) return false;
        c.c_out = (int)sh[0];
        c.c_in = (int)sh[1];
        c.k = (int)sh[2];
        c.dilation = dilation;
        c.stride = 1;
        c.pad = ((c.k - 1) * dilation) / 2;
        if (c.c_in % 8 != 0) return fail(4, "unsupported channel count (not a multiple of 8) in " + prefix);
        std::vector<float> r((size_t)c.c_out * c.k * c.c_in);
        for (int co = 0; co < c.c_out; ++co)
            for (int ci = 0; ci < c.c_in; ++ci)
                for (int t = 0; t < c.k; ++t) r[((size_t)co * c.k + t) * c.c_in + ci] = w[((size_t)co * c.c_in + ci) * c.k + t];
        std::vector<uint8_t> img;
        std::string e;
        if (!pack_tiles_f32(r.data(), c.c_out, c.k, c.c_in, W_F16, img, c.w, e)) return fail(1, e);
        c.w.dev = dev_upload(img.data(), img.size());
        if (!c.w.dev) return false;
        if (with_bias && !bias_vec(prefix + ".bias", c.c_out, 1, c.bias)) return false;
        return true;
    }
    // conv-transpose (k = 2*stride): weight_v [Cin, Cout, K] -> W'[(phase r, co)][tap][ci] = w[ci][co][r + tap*stride]
    bool conv_t(const std::string& prefix, int stride, VaeConv& c) {
        std::vector<float> w;
        std::vector<int64_t> sh;
        if (!folded(prefix, w, sh)) return false;
        c.c_in = (int)sh[0];
        c.c_out = (int)sh[1];
        c.k = (int)sh[2];
        c.stride = stride;
        c.dilation = 1;
        c.pad = (stride + 1) / 2;
        if (c.k != 2 * stride) return fail(4, "unsupported conv-transpose kernel size (expected 2*stride) in " + prefix);
        if (c.c_in % 8 != 0 || c.c_out % 32 != 0) return fail(4, "unsupported channel count in " + prefix);
        const int N = stride * c.c_out;
        std::vector<float> r((size_t)N * 2 * c.c_in);
        for (int ph = 0; ph < stride; ++ph)
            for (int co = 0; co < c.c_out; ++co)
                for (int tap = 0; tap < 2; ++tap)
                    for (int ci = 0; ci < c.c_in; ++ci)
                        r[(((size_t)ph * c.c_out + co) * 2 + tap) * c.c_in + ci] = w[((size_t)ci * c.c_out + co) * c.k + ph + tap * stride];
        std::vector<uint8_t> img;
        std::string e;
        if (!pack_tiles_f32(r.data(), N, 2, c.c_in, W_F16, img, c.w, e)) return fail(1, e);
        c.w.dev = dev_upload(img.data(), img.size());
        if (!c.w.dev) return false;
        return bias_vec(prefix + ".bias", c.c_out, stride, c.bias);
    }
    // plain conv whose input channels are zero-padded to c_pad (encoder.conv1: 2 audio channels -> 8, the GEMM's row granule)
    bool conv_padded(const std::string& prefix, int c_pad, VaeConv& c) {
        std::vector<float> w;
        std::vector<int64_t> sh;
        if (!folded(prefix, w, sh)) return false;
        c.c_out = (int)sh[0];
        const int cin = (int)sh[1];
        c.c_in = c_pad;
        c.k = (int)sh[2];
        c.dilation = 1;
        c.stride = 1;
        c.pad = (c.k - 1) / 2;
        if (cin > c_pad) return fail(4, "unsupported channel count in " + prefix);
        std::vector<float> r((size_t)c.c_out * c.k * c_pad, 0.f);
        for (int co = 0; co < c.c_out; ++co)
            for (int ci = 0; ci < cin; ++ci)
                for (int t = 0; t < c.k; ++t) r[((size_t)co * c.k + t) * c_pad + ci] = w[((size_t)co * cin + ci) * c.k + t];
        std::vector<uint8_t> img;
        std::string e;
        if (!pack_tiles_f32(r.data(), c.c_out, c.k, c_pad, W_F16, img, c.w, e)) return fail(1, e);
        c.w.dev = dev_upload(img.data(), img.size());
        return c.w.dev && bias_vec(prefix + ".bias", c.c_out, 1, c.bias);
    }
    // strided conv (k = 2*stride, stride s, pad ceil(s/2)): weight_v [Cout, Cin, K].  With the padded input viewed as rows of s
    // consecutive positions ([L', s*Cin]) it is a 2-tap GEMM: W'[co][T][j*Cin + ci] = w[co][ci][T*s + j]
    bool conv_strided(const std::string& prefix, int stride, VaeConv& c) {
        std::vector<float> w;
        std::vector<int64_t> sh;
        if (!folded(prefix, w, sh)) return false;
        c.c_out = (int)sh[0];
        c.c_in = (int)sh[1];
        c.k = (int)sh[2];
        c.stride = stride;
        c.dilation = 1;
        c.pad = (stride + 1) / 2;
        if (c.k != 2 * stride) return fail(4, "unsupported strided-conv kernel size (expected 2*stride) in " + prefix);
        if (c.c_in % 8 != 0) return fail(4, "unsupported channel count in " + prefix);
        const int K = stride * c.c_in;
        std::vector<float> r((size_t)c.c_out * 2 * K);
        for (int co = 0; co < c.c_out; ++co)
            for (int T = 0; T < 2; ++T)
                for (int j = 0; j < stride; ++j)
                    for (int ci = 0; ci < c.c_in; ++ci)
                        r[((size_t)co * 2 + T) * K + (size_t)j * c.c_in + ci] = w[((size_t)co * c.c_in + ci) * c.k + T * stride + j];
        std::vector<uint8_t> img;
        std::string e;
        if (!pack_tiles_f32(r.data(), c.c_out, 2, K, W_F16, img, c.w, e)) return fail(1, e);
        c.w.dev = dev_upload(img.data(), img.size());
        return c.w.dev && bias_vec(prefix + ".bias", c.c_out, 1, c.bias);
    }
    bool res_unit(const std::string& prefix, int C, int dilation, VaeResUnit& ru) {
        return snake(prefix + ".snake1", C, ru.snake1) && conv(prefix + ".conv1", dilation, true, ru.conv1) &&
               snake(prefix + ".snake2", C, ru.snake2) && conv(prefix + ".conv2", 1, true, ru.conv2);
    }
};

bool parse_vae_config(const std::string& path, VaeConfig& c, std::string& err) {
    JsonValue root;
    if (!json_parse_file(path, root, err)) return false;
    if (root.kind != JsonValue::Object) { err = "VAE config is not a JSON object"; return false; }
    auto gi = [&](const char* k, int& dst) {
        int64_t v;
        if (!root.get_int(k, v)) { err = std::string("missing or invalid integer key: ") + k; return false; }
        dst = (int)v;
        return true;
    };
    auto ga = [&](const char* k, std::vector<int>& dst) {
        const JsonValue* a = root.get(k);
        if (!a || a->kind != JsonValue::Array) { err = std::string("missing or invalid array key: ") + k; return false; }
        dst.clear();
        for (const auto& v : a->arr) dst.push_back((int)v.num);
        return true;
    };
    if (!gi("audio_channels", c.audio_channels) || !gi("encoder_hidden_size", c.encoder_hidden_size) ||
        !gi("decoder_channels", c.decoder_channels) || !gi("decoder_input_channels", c.decoder_input_channels) ||
        !gi("sampling_rate", c.sampling_rate) || !ga("downsampling_ratios", c.downsampling_ratios) ||
        !ga("channel_multiples", c.channel_multiples))
        return false;
    if (c.downsampling_ratios.empty() || c.channel_multiples.empty()) { err = "invalid VAE config: empty ratios or channel multiples"; return false; }
    c.upsampling_ratios.assign(c.downsampling_ratios.rbegin(), c.downsampling_ratios.rend());
    c.hop_length = 1;
    for (int r : c.downsampling_ratios) c.hop_length *= r;
    return true;
}

}  // namespace

int vae_load(const std::string& model_dir_in, VaeModel& m, std::string& err) {
    m.release();
    m.cfg = VaeConfig();
    // accept <dir> or <dir>/vae (acestep_ggml.cpp:534-555)
    std::string dir = model_dir_in;
    if (!file_exists(dir + "/config.json") && file_exists(dir + "/vae/config.json")) dir += "/vae";
    if (!parse_vae_config(dir + "/config.json", m.cfg, err)) return 3;
    const VaeConfig& c = m.cfg;

    std::unique_ptr<TensorFile> tf;
    std::string gguf;
    for (const char* key : {"ACE_GGML_VAE_GGUF", "ACE_GGML_VAE_GGUF_PATH"}) {
        const char* v = std::getenv(key);
        if (gguf.empty() && v && v[0] && file_exists(v)) gguf = v;
    }
    if (gguf.empty() && file_exists(dir + "/model.gguf")) gguf = dir + "/model.gguf";
    if (!gguf.empty()) tf = TensorFile::open_gguf(gguf, err);
    else tf = TensorFile::open_safetensors(dir + "/diffusion_pytorch_model.safetensors", err);
    if (!tf) return 3;

    VLoader L{tf.get(), &m, {}, 0};
    // channel plan: decoder_channels * reversed([1] + multiples)
    std::vector<int> mult;
    mult.push_back(1);
    for (int x : c.channel_multiples) mult.push_back(x);
    std::vector<int> ch;
    for (auto it = mult.rbegin(); it != mult.rend(); ++it) ch.push_back(c.decoder_channels * *it);
    const int nb = (int)c.upsampling_ratios.size();
    if ((int)ch.size() < nb + 1) { err = "invalid VAE config: channel_multiples shorter than ratios"; return 3; }

    bool ok = L.conv("decoder.conv1", 1, true, m.conv1);
    if (ok && (m.conv1.c_in != c.decoder_input_channels || m.conv1.c_out != ch[0] || ch[0] % 32 != 0)) ok = L.fail(3, "unexpected decoder.conv1 shape");
    m.blocks.resize((size_t)nb);
    static const int dil[3] = {1, 3, 9};
    for (int i = 0; ok && i < nb; ++i) {
        VaeBlock& b = m.blocks[(size_t)i];
        const std::string p = "decoder.block." + std::to_string(i);
        ok = ok && L.snake(p + ".snake1", ch[(size_t)i], b.snake1);
        ok = ok && L.conv_t(p + ".conv_t1", c.upsampling_ratios[(size_t)i], b.conv_t);
        if (ok && (b.conv_t.c_in != ch[(size_t)i] || b.conv_t.c_out != ch[(size_t)i + 1])) ok = L.fail(3, "unexpected shape for " + p + ".conv_t1");
        for (int j = 0; ok && j < 3; ++j) {
            ok = L.res_unit(p + ".res_unit" + std::to_string(j + 1), ch[(size_t)i + 1], dil[j], b.res[j]);
            if (ok && (b.res[j].conv1.c_out != ch[(size_t)i + 1] || b.res[j].conv1.k != 7 || b.res[j].conv2.k != 1)) ok = L.fail(3, "unexpected residual-unit shape in " + p);
        }
    }
    ok = ok && L.snake("decoder.snake1", ch[(size_t)nb], m.snake_out);
    if (ok) {  // decoder.conv2: k7, no bias, audio_channels outputs -> fp16 [Cout][tap][ci] for the tail kernel
        std::vector<float> w;
        std::vector<int64_t> sh;
        if (L.folded("decoder.conv2", w, sh)) {
            const int co = (int)sh[0], ci = (int)sh[1], k = (int)sh[2];
            if (co != c.audio_channels || ci != ch[(size_t)nb] || k != 7 || co > 2 || ci % 8 != 0 || ci > 512) ok = L.fail(4, "unsupported decoder.conv2 shape");
            else {
                std::vector<uint16_t> h((size_t)co * k * ci);
                for (int o = 0; o < co; ++o)
                    for (int i2 = 0; i2 < ci; ++i2)
                        for (int t = 0; t < k; ++t) h[((size_t)o * k + t) * ci + i2] = f32_to_f16(w[((size_t)o * ci + i2) * k + t]);
                m.conv2_w = static_cast<__half*>(L.dev_upload(h.data(), h.size() * 2));
                m.conv2_cin = ci;
                ok = m.conv2_w != nullptr;
            }
        } else ok = false;
    }
    // ---- encoder (acestep_vae_model.cpp:913-929): conv1 k7 -> blocks [3 res units, snake, strided conv] -> snake -> conv2 k3
    if (ok && tf->find("encoder.conv1.weight_v")) {
        std::vector<int> ech;  // encoder_hidden_size * ([1] + multiples)
        for (int x : mult) ech.push_back(c.encoder_hidden_size * x);
        const int ne = (int)c.downsampling_ratios.size();
        ok = (int)ech.size() >= ne + 1 || L.fail(3, "invalid VAE config: channel_multiples shorter than ratios");
        ok = ok && L.conv_padded("encoder.conv1", 8, m.enc_conv1);
        if (ok && (m.enc_conv1.c_out != ech[0] || m.enc_conv1.k != 7)) ok = L.fail(3, "unexpected encoder.conv1 shape");
        m.enc_blocks.resize((size_t)ne);
        for (int i = 0; ok && i < ne; ++i) {
            VaeEncBlock& b = m.enc_blocks[(size_t)i];
            const std::string p = "encoder.block." + std::to_string(i);
            for (int j = 0; ok && j < 3; ++j) {
                ok = L.res_unit(p + ".res_unit" + std::to_string(j + 1), ech[(size_t)i], dil[j], b.res[j]);
                if (ok && (b.res[j].conv1.c_out != ech[(size_t)i] || b.res[j].conv1.k != 7 || b.res[j].conv2.k != 1)) ok = L.fail(3, "unexpected residual-unit shape in " + p);
            }
            ok = ok && L.snake(p + ".snake1", ech[(size_t)i], b.snake1);
            ok = ok && L.conv_strided(p + ".conv1", c.downsampling_ratios[(size_t)i], b.conv);
            if (ok && (b.conv.c_in != ech[(size_t)i] || b.conv.c_out != ech[(size_t)i + 1])) ok = L.fail(3, "unexpected shape for " + p + ".conv1");
        }
        ok = ok && L.snake("encoder.snake1", ech[(size_t)ne], m.enc_snake);
        ok = ok && L.conv("encoder.conv2", 1, true, m.enc_conv2);
        if (ok && (m.enc_conv2.c_in != ech[(size_t)ne] || m.enc_conv2.c_out < c.decoder_input_channels || m.enc_conv2.k != 3))
            ok = L.fail(3, "unexpected encoder.conv2 shape");
        m.has_encoder = ok;
    }
    if (!ok) {
        err = L.err;
        m.release();
        return L.status ? L.status : 1;
    }
    return 0;
}

// ------------------------------------------------------------------------------------------------ encoder
long vae_encode_out_frames(const VaeModel& m, long n_samples) {
    long L = n_samples;
    for (const VaeEncBlock& b : m.enc_blocks) {
        const long s = b.conv.stride, p = b.conv.pad;
        if (L + 2 * p < 2 * s) return 0;
        L = (L + 2 * p - 2 * s) / s + 1;
    }
    return L;
}

bool vae_encode_workspace_reserve(const VaeModel& m, VaeEncWorkspace& ws, long n_samples, std::string& err) {
    if ((size_t)n_samples <= ws.cap_samples && !ws.allocs.empty()) return true;
    const size_t S = std::max<size_t>((size_t)n_samples, ws.cap_samples);
    ws.release();
    // widest level in elements: channels x positions never grows along the encoder (channels x2, positions / stride >= 2)
    size_t lvl = (size_t)m.enc_conv1.c_out * S, widest = lvl, L = S;
    for (const VaeEncBlock& b : m.enc_blocks) {
        widest = std::max(widest, (size_t)b.conv.c_in * (L + 4 * (size_t)b.conv.stride));  // padded copy of the level
        L = L / (size_t)b.conv.stride + 2;
        widest = std::max(widest, (size_t)b.conv.c_out * L);
    }
    bool ok = true;
    auto alloc = [&](auto*& ptr, size_t count) {
        if (!ok) return;
        void* p = nullptr;
        if (cudaMalloc(&p, count * sizeof(*ptr) + 256) != cudaSuccess) {
            err = std::string("VAE encoder workspace cudaMalloc failed: ") + cudaGetErrorString(cudaGetLastError());
            ok = false;
            return;
        }
        ws.allocs.push_back(p);
        ptr = reinterpret_cast<decltype(ptr)>(p);
    };
    alloc(ws.a16, S * 8);
    alloc(ws.x, widest);
    alloc(ws.xs, widest);
    alloc(ws.h, widest);
    alloc(ws.xp, widest);
    alloc(ws.y, (L + 2) * (size_t)m.enc_conv2.c_out);
    if (!ok) {
        ws.release();
        return false;
    }
    ws.cap_samples = S;
    return true;
}

cudaError_t vae_encode_device(const VaeModel& m, VaeEncWorkspace& ws, const float* d_audio, long n_samples, float* d_latents,
                              cudaStream_t st) {
    const VaeConfig& c = m.cfg;
    cudaError_t rc = cudaSuccess;
    auto run = [&](const __half* A, long L, long l_out, int C, const VaeConv& cv, int tap_stride, int tap_offset, GemmEpilogue e) {
        if (rc != cudaSuccess) return;
        GemmProblem g;
        g.A = A;
        g.B = 1;
        g.L = (int)L;
        g.C = C;
        g.a_lstride = C;
        g.a_bstride = L * C;
        g.l_out = (int)l_out;
        g.tap_stride = tap_stride;
        g.tap_offset = tap_offset;
        g.W = &cv.w;
        g.epi = e;
        if (g.epi.out_bstride == 0) g.epi.out_bstride = l_out * g.epi.ld_out;
        rc = gemm_launch(g, st);
    };
    auto with_snake = [](GemmEpilogue e, const VaeSnake& s) {
        e.snake_alpha = s.alpha;
        e.snake_invbeta = s.inv_beta;
        return e;
    };
    long L = n_samples;
    launch_pad_channels_f16(d_audio, ws.a16, L, c.audio_channels, 8, st);
    const int ne = (int)m.enc_blocks.size();
    {  // conv1 k7: epilogue = bias -> x (fp32) and block0.res1.snake1 -> xs
        GemmEpilogue e;
        e.bias = m.enc_conv1.bias;
        e.out_f32 = ws.x;
        e.out_f16 = ws.xs;
        e.ld_out = m.enc_conv1.c_out;
        run(ws.a16, L, L, 8, m.enc_conv1, 1, -m.enc_conv1.pad, with_snake(e, ne ? m.enc_blocks[0].res[0].snake1 : m.enc_snake));
    }
    for (int i = 0; i < ne; ++i) {
        const VaeEncBlock& b = m.enc_blocks[(size_t)i];
        const int C = b.conv.c_in, s = b.conv.stride, p = b.conv.pad;
        for (int j = 0; j < 3; ++j) {
            const VaeResUnit& ru = b.res[j];
            {
                GemmEpilogue e;
                e.bias = ru.conv1.bias;
                e.out_f16 = ws.h;
                e.ld_out = C;
                run(ws.xs, L, L, C, ru.conv1, ru.conv1.dilation, -ru.conv1.pad, with_snake(e, ru.snake2));
            }
            {
                GemmEpilogue e;
                e.bias = ru.conv2.bias;
                e.resid = ws.x;
                e.ld_out = C;
                e.out_f16 = ws.xs;
                if (j < 2) e.out_f32 = ws.x;
                run(ws.h, L, L, C, ru.conv2, 1, 0, with_snake(e, j < 2 ? b.res[j + 1].snake1 : b.snake1));
            }
        }
        // strided conv: xp = [p zero rows | snake(x) | zero tail], read as rows of s positions; out[l] = xp'[l] W0 + xp'[l+1] W1
        const long Lout = (L + 2 * p - 2 * s) / s + 1;
        const long rows = Lout + 1;
        if (rc == cudaSuccess) rc = cudaMemsetAsync(ws.xp, 0, (size_t)p * C * 2, st);
        if (rc == cudaSuccess) rc = cudaMemcpyAsync(ws.xp + (size_t)p * C, ws.xs, (size_t)L * C * 2, cudaMemcpyDeviceToDevice, st);
        const long used = (long)p + L, total = rows * s;
        if (rc == cudaSuccess && total > used) rc = cudaMemsetAsync(ws.xp + (size_t)used * C, 0, (size_t)(total - used) * C * 2, st);
        {
            GemmEpilogue e;
            e.bias = b.conv.bias;
            e.ld_out = b.conv.c_out;
            e.out_f16 = ws.xs;
            if (i + 1 < ne) e.out_f32 = ws.x;
            run(ws.xp, rows, Lout, s * C, b.conv, 1, 0, with_snake(e, i + 1 < ne ? m.enc_blocks[(size_t)i + 1].res[0].snake1 : m.enc_snake));
        }
        L = Lout;
    }
    {  // conv2 k3 pad 1 -> [L, encoder_hidden_size]; the posterior mean is the first latent_ch channels (:1039-1043)
        GemmEpilogue e;
        e.bias = m.enc_conv2.bias;
        e.out_f32 = ws.y;
        e.ld_out = m.enc_conv2.c_out;
        run(ws.xs, L, L, m.enc_conv2.c_in, m.enc_conv2, 1, -m.enc_conv2.pad, e);
    }
    if (rc != cudaSuccess) return rc;
    launch_take_channels_f32(ws.y, d_latents, L, m.enc_conv2.c_out, c.decoder_input_channels, st);
    return cudaGetLastError();
}

bool vae_workspace_reserve(const VaeModel& m, VaeWorkspace& ws, int B, int n_frames, std::string& err) {
    if ((size_t)n_frames <= ws.cap_frames && B <= ws.capB && !ws.allocs.empty()) return true;
    const size_t F = std::max<size_t>((size_t)n_frames, ws.cap_frames);
    const int nB = std::max(B, ws.capB);
    ws.release();
    // per-frame element counts at every level
    size_t max_level = (size_t)m.conv1.c_out, max_in = (size_t)m.conv1.c_out;
    size_t L = 1;
    for (size_t i = 0; i < m.blocks.size(); ++i) {
        L *= (size_t)m.blocks[i].conv_t.stride;
        const size_t lvl = L * (size_t)m.blocks[i].conv_t.c_out;
        max_level = std::max(max_level, lvl);
        if (i + 1 < m.blocks.size()) max_in = std::max(max_in, lvl);
    }
    bool ok = true;
    auto alloc = [&](auto*& ptr, size_t count) {
        if (!ok) return;
        void* p = nullptr;
        if (cudaMalloc(&p, count * sizeof(*ptr) + 256) != cudaSuccess) {
            err = std::string("VAE workspace cudaMalloc failed: ") + cudaGetErrorString(cudaGetLastError());
            ok = false;
            return;
        }
        ws.allocs.push_back(p);
        ptr = reinterpret_cast<decltype(ptr)>(p);
    };
    const size_t BF = (size_t)nB * F;
    alloc(ws.d_lat, BF * m.cfg.decoder_input_channels);
    alloc(ws.lat16, BF * m.cfg.decoder_input_channels);
    alloc(ws.x, BF * max_level);
    alloc(ws.xs, BF * max_level);
    alloc(ws.h, BF * max_level);
    // block inputs ping-pong with xs: reuse a region that is never live at the same time is not possible for every
    // level, so they get their own buffer (<= half of a full level)
    __half* blk = nullptr;
    alloc(blk, BF * max_in);
    if (!ok) {
        ws.release();
        return false;
    }
    ws.cap_frames = F;
    ws.capB = nB;
    return true;
}

cudaError_t vae_decode_device(const VaeModel& m, VaeWorkspace& ws, int B, int n_frames, float* audio, cudaStream_t st,
                              const VaeOutWindow* win) {
    const VaeConfig& c = m.cfg;
    __half* blk_in = static_cast<__half*>(ws.allocs.back());
    cudaError_t rc = cudaSuccess;
    auto run = [&](const __half* A, int L, int l_out, const VaeConv& cv, int tap_stride, int tap_offset, GemmEpilogue e) {
        if (rc != cudaSuccess) return;
        GemmProblem g;
        g.A = A;
        g.B = B;
        g.L = L;
        g.C = cv.c_in;
        g.a_lstride = cv.c_in;
        g.a_bstride = (long)L * cv.c_in;
        g.l_out = l_out;
        g.tap_stride = tap_stride;
        g.tap_offset = tap_offset;
        g.W = &cv.w;
        g.epi = e;
        rc = gemm_launch(g, st);
    };
    auto with_snake = [](GemmEpilogue e, const VaeSnake& s, int period) {
        e.snake_alpha = s.alpha;
        e.snake_invbeta = s.inv_beta;
        e.snake_period = period;
        return e;
    };

    // Residual stream between the convolutions: 16-bit by default (the k1 convolutions of the last levels are HBM-bound on it:
    // fp32 in + fp32 out + fp16 Snake'd copy = 10 of their 12 bytes per element; fp16 in / out makes it 8), fp32 with
    // ACE_B200_VAE_X32=1 (round 1's layout).  The reference adds in F32 and rounds every conv INPUT to F16 (ggml im2col), so the extra
    // rounding sits on values that are rounded one op later anyway; the waveform bound (cos >= 0.9999, SNR >= 40 dB) holds with margin.
    const char* x32_env = std::getenv("ACE_B200_VAE_X32");
    const bool x16 = !(x32_env && x32_env[0] && x32_env[0] != '0');
    __half* xh = reinterpret_cast<__half*>(ws.x);
    launch_f32_to_f16(ws.d_lat, ws.lat16, (long)B * n_frames * c.decoder_input_channels, st);
    long L = n_frames;
    {  // conv1 k7: latents -> C0 channels, epilogue = bias + block0.snake1
        GemmEpilogue e;
        e.bias = m.conv1.bias;
        e.out_f16 = blk_in;
        e.ld_out = m.conv1.c_out;
        e.out_bstride = L * m.conv1.c_out;
        run(ws.lat16, (int)L, (int)L, m.conv1, 1, -m.conv1.pad, with_snake(e, m.blocks[0].snake1, 0));
    }
    const int nb = (int)m.blocks.size();
    const char* fuse_env = std::getenv("ACE_B200_VAE_FUSE");   // 0 = every unit as two GEMM launches, 1 = fused on single CTAs (A/B and bisecting aids)
    const bool fuse_units = !(fuse_env && fuse_env[0] == '0');
    const __half* tail_in = ws.xs;  // input of the final k7 convolution
    for (int i = 0; i < nb; ++i) {
        const VaeBlock& b = m.blocks[(size_t)i];
        const int s = b.conv_t.stride, Co = b.conv_t.c_out, p = b.conv_t.pad;
        const long Lout = L * s + s - 2L * p;
        {  // conv-transpose as `s` phase-GEMMs in one launch: D[i, (r, co)] = out_full[i*s + r, co]; crop = flat shift
            GemmEpilogue e;
            e.bias = b.conv_t.bias;
            if (x16) e.out_x16 = xh;
            else e.out_f32 = ws.x;
            e.out_f16 = ws.xs;
            e.ld_out = (long)s * Co;
            e.out_bstride = Lout * Co;
            e.out_shift = -(long)p * Co;
            e.out_valid = Lout * Co;
            run(blk_in, (int)L, (int)L + 1, b.conv_t, -1, 0, with_snake(e, b.res[0].snake1, Co));
        }
        L = Lout;
        // Fused residual units (vae_resunit.cu) at 128 channels with the 16-bit residual stream: the Snake'd copy ping-pongs between
        // ws.xs and ws.h (a tile's k7 taps read its neighbours' rows, so a unit cannot overwrite its own input).
        const bool fuse = fuse_units && x16 && vae_resunit128_supported(b.res[0]) && vae_resunit128_supported(b.res[1]) &&
                          vae_resunit128_supported(b.res[2]);
        for (int j = 0; j < 3; ++j) {
            const VaeResUnit& ru = b.res[j];
            if (fuse) {
                if (rc != cudaSuccess) break;
                const __half* in = (j & 1) ? ws.h : ws.xs;
                __half* out = (j & 1) ? ws.xs : ws.h;
                const VaeSnake* next = j < 2 ? &b.res[j + 1].snake1 : (i + 1 < nb ? &m.blocks[(size_t)i + 1].snake1 : &m.snake_out);
                if (j == 2 && i + 1 < nb) out = blk_in;
                rc = launch_vae_resunit128(in, xh, j < 2 ? xh : nullptr, out, B, L, ru, *next, st);
                if (j == 2 && i + 1 == nb) tail_in = out;
                continue;
            }
            {  // k7 dilated conv, epilogue = bias + snake2 -> h
                GemmEpilogue e;
                e.bias = ru.conv1.bias;
                e.out_f16 = ws.h;
                e.ld_out = Co;
                e.out_bstride = L * Co;
                run(ws.xs, (int)L, (int)L, ru.conv1, ru.conv1.dilation, -ru.conv1.pad, with_snake(e, ru.snake2, 0));
            }
            {  // k1 conv, epilogue = bias + skip, then the NEXT consumer's snake
                GemmEpilogue e;
                e.bias = ru.conv2.bias;
                if (x16) e.resid16 = xh;
                else e.resid = ws.x;
                e.ld_out = Co;
                e.out_bstride = L * Co;
                const VaeSnake* next;
                if (j < 2) {
                    if (x16) e.out_x16 = xh;
                    else e.out_f32 = ws.x;
                    e.out_f16 = ws.xs;
                    next = &b.res[j + 1].snake1;
                } else if (i + 1 < nb) {
                    e.out_f16 = blk_in;
                    next = &m.blocks[(size_t)i + 1].snake1;
                } else {
                    e.out_f16 = ws.xs;
                    next = &m.snake_out;
                }
                run(ws.h, (int)L, (int)L, ru.conv2, 1, 0, with_snake(e, *next, 0));
            }
        }
    }
    if (rc != cudaSuccess) return rc;
    if (win) {
        const long hop = c.hop_length;
        launch_conv_small_out(tail_in, m.conv2_w, nullptr, audio, B, L, m.conv2_cin, c.audio_channels, 7, 3, st, win->frame_lo * hop,
                              win->frame_hi * hop, (win->dst_frame - win->frame_lo) * hop, 0);
    } else {
        launch_conv_small_out(tail_in, m.conv2_w, nullptr, audio, B, L, m.conv2_cin, c.audio_channels, 7, 3, st);
    }
    return cudaGetLastError();
}

}  // namespace ace
