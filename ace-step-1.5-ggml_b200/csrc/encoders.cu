// Transformer stack shared by the lyric / timbre condition encoders and the Qwen3 text encoder (see encoders.h).
#include "encoders.h"
#include "kernels.h"

#include <algorithm>

namespace ace {

bool enc_stack_load(WeightLoader& L, const std::string& prefix, int n_layers, EncStack& st) {
    const int H = st.H, I = st.I, nq = st.n_heads * st.hd, nkv = st.n_kv * st.hd;
    st.layers.resize((size_t)n_layers);
    for (int i = 0; i < n_layers; ++i) {
        EncLayer& ly = st.layers[(size_t)i];
        const std::string p = prefix + std::to_string(i) + ".";
        bool ok = L.vec(p + "input_layernorm.weight", H, ly.input_norm) && L.vec(p + "post_attention_layernorm.weight", H, ly.post_norm) &&
                  L.linear({p + "self_attn.q_proj.weight", p + "self_attn.k_proj.weight", p + "self_attn.v_proj.weight"}, false, H, ly.qkv) &&
                  L.linear({p + "self_attn.o_proj.weight"}, false, nq, ly.o) && L.vec(p + "self_attn.q_norm.weight", st.hd, ly.qn) &&
                  L.vec(p + "self_attn.k_norm.weight", st.hd, ly.kn) &&
                  L.linear({p + "mlp.gate_proj.weight", p + "mlp.up_proj.weight"}, true, H, ly.gate_up) &&
                  L.linear({p + "mlp.down_proj.weight"}, false, I, ly.down);
        if (!ok) return false;
        if (ly.qkv.N != nq + 2 * nkv || ly.o.N != H || ly.gate_up.N != 2 * I || ly.down.N != H)
            return L.fail(3, "unexpected projection shape in " + p);
    }
    return true;
}

void EncWorkspace::release() {
    for (void* p : allocs) cudaFree(p);
    allocs.clear();
    *this = EncWorkspace();
}

bool enc_workspace_reserve(const EncStack& st, EncWorkspace& ws, int rows, std::string& err) {
    const size_t nq = (size_t)st.n_heads * st.hd, nkv = (size_t)st.n_kv * st.hd;
    const size_t widest = std::max<size_t>(std::max<size_t>((size_t)st.H, (size_t)st.I), nq + 2 * nkv);
    const size_t need = (size_t)rows * widest;
    if (!ws.allocs.empty() && rows <= ws.cap_rows && need <= ws.cap_elems) return true;
    const int R = std::max(rows, ws.cap_rows);
    const size_t elems = std::max(need, ws.cap_elems);
    ws.release();
    bool ok = true;
    auto alloc = [&](auto*& ptr, size_t count) {
        if (!ok) return;
        void* p = nullptr;
        if (cudaMalloc(&p, count * sizeof(*ptr) + 256) != cudaSuccess) {
            err = std::string("encoder workspace cudaMalloc failed: ") + cudaGetErrorString(cudaGetLastError());
            ok = false;
            return;
        }
        ws.allocs.push_back(p);
        ptr = reinterpret_cast<decltype(ptr)>(p);
    };
    // every buffer is sized for the widest row so that one workspace serves stacks of different widths (lyric / timbre / text)
    alloc(ws.x, elems);
    alloc(ws.xn, elems);
    alloc(ws.qkv, elems);
    alloc(ws.att, elems);
    alloc(ws.act, elems);
    alloc(ws.rope, (size_t)R * 64);
    alloc(ws.d_mask, (size_t)R);
    if (!ok) {
        ws.release();
        return false;
    }
    ws.cap_rows = R;
    ws.cap_elems = elems;
    ws.rope_len = 0;
    return true;
}

cudaError_t enc_stack_forward(const EncStack& st, EncWorkspace& ws, int rows, int n_layers, bool causal, const int32_t* key_mask,
                              bool final_norm, float* out, cudaStream_t stream) {
    const int H = st.H, I = st.I, hd = st.hd;
    const int nq = st.n_heads * hd, nkv = st.n_kv * hd, qkv_w = nq + 2 * nkv;
    const int total = (int)st.layers.size();
    const int run = (n_layers < 0 || n_layers > total) ? total : n_layers;
    if (ws.rope_len < rows || ws.rope_theta != st.theta) {
        launch_build_rope_table(ws.rope, ws.cap_rows, hd / 2, st.theta, stream);
        ws.rope_len = ws.cap_rows;
        ws.rope_theta = st.theta;
    }
    cudaError_t rc = cudaSuccess;
    auto gemm = [&](const __half* A, long lda, const PackedWeight& W, GemmEpilogue epi, bool accumulate) {
        if (rc != cudaSuccess) return;
        GemmProblem g;
        g.A = A;
        g.B = 1;
        g.L = rows;
        g.C = W.K;
        g.a_lstride = lda;
        g.a_bstride = (long)rows * lda;
        g.l_out = rows;
        g.W = &W;
        g.epi = epi;
        if (g.epi.out_bstride == 0) g.epi.out_bstride = (long)rows * g.epi.ld_out;
        g.split_k = accumulate ? 0 : 1;
        rc = gemm_launch(g, stream);
    };
    for (int i = 0; i < run; ++i) {
        const EncLayer& ly = st.layers[(size_t)i];
        launch_rmsnorm_mod(ws.x, ly.input_norm, nullptr, nullptr, 0, ws.xn, 1, rows, H, st.eps, stream);
        {
            GemmEpilogue e;  // per-head RMSNorm(q, k) + NEOX RoPE fused (positions 0..rows-1), V stored as is
            e.mode = EPI_HEADNORM;
            e.hn_wa = ly.qn;
            e.hn_na = nq;
            e.hn_wb = ly.kn;
            e.hn_nb = nkv;
            e.rope = ws.rope;
            e.hn_eps = st.eps;
            e.out_f16 = ws.qkv;
            e.ld_out = qkv_w;
            gemm(ws.xn, H, ly.qkv, e, false);
        }
        launch_attention(ws.qkv, qkv_w, ws.qkv + nq, qkv_w, ws.qkv + nq + nkv, qkv_w, ws.att, nq, key_mask, 1, rows, rows, st.n_heads,
                         st.n_kv, ly.sliding ? st.sliding_window : -1, stream, causal);
        {
            GemmEpilogue e;
            e.resid = ws.x;
            e.out_f32 = ws.x;
            e.ld_out = H;
            gemm(ws.att, nq, ly.o, e, true);
        }
        launch_rmsnorm_mod(ws.x, ly.post_norm, nullptr, nullptr, 0, ws.xn, 1, rows, H, st.eps, stream);
        {
            GemmEpilogue e;
            e.mode = EPI_SWIGLU;
            e.out_f16 = ws.act;
            e.ld_out = I;
            gemm(ws.xn, H, ly.gate_up, e, false);
        }
        {
            GemmEpilogue e;
            e.resid = ws.x;
            e.out_f32 = ws.x;
            e.ld_out = H;
            gemm(ws.act, I, ly.down, e, true);
        }
    }
    if (rc != cudaSuccess) return rc;
    if (final_norm && st.final_norm) {
        launch_rmsnorm_mod(ws.x, st.final_norm, nullptr, nullptr, 0, nullptr, 1, rows, H, st.eps, stream, out);
    } else {
        rc = cudaMemcpyAsync(out, ws.x, (size_t)rows * H * 4, cudaMemcpyDeviceToDevice, stream);
        if (rc != cudaSuccess) return rc;
    }
    return cudaGetLastError();
}

}  // namespace ace
