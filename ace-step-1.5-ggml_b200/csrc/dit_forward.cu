// Batched DiT forward on the device: a static launch schedule of tap-GEMMs (tcgen05), fused attention and
// element-wise kernels.   reference graph: acestep_dit_model.cpp:1316-1560 (forward_dit), 1175-1259 (attention),
// 1286-1308 (timestep_forward)
#include "dit.h"
#include "gemm.h"
#include "kernels.h"

#include <algorithm>

namespace ace {

void DitWorkspace::release() {
    for (void* p : allocs) cudaFree(p);
    allocs.clear();
    *this = DitWorkspace();
}

bool dit_workspace_reserve(const DitModel& m, DitWorkspace& ws, int B, int T, int E, std::string& err) {
    if (B <= ws.capB && T <= ws.capT && E <= ws.capE && !ws.allocs.empty()) return true;
    const int nB = std::max(B, ws.capB), nT = std::max(T, ws.capT), nE = std::max(std::max(E, ws.capE), 1);
    ws.release();
    const DitConfig& c = m.cfg;
    const int P = c.patch_size, H = c.hidden_size, I = c.intermediate_size, A = c.audio_acoustic_hidden_dim;
    const int ctx = c.in_channels - A;
    const long S = (nT + P - 1) / P;
    const long nq = (long)c.num_attention_heads * c.head_dim, nkv = (long)c.num_key_value_heads * c.head_dim;
    const long rows = (long)nB * S, erows = (long)nB * nE;
    const int layers = c.num_hidden_layers;
    bool ok = true;
    auto alloc = [&](auto*& ptr, size_t count) {
        if (!ok) return;
        void* p = nullptr;
        const size_t bytes = count * sizeof(*ptr);
        if (cudaMalloc(&p, bytes ? bytes : 16) != cudaSuccess) {
            err = std::string("workspace cudaMalloc failed: ") + cudaGetErrorString(cudaGetLastError());
            ok = false;
            return;
        }
        ws.allocs.push_back(p);
        ptr = reinterpret_cast<decltype(ptr)>(p);
    };
    alloc(ws.d_hidden, (size_t)nB * nT * A);
    alloc(ws.d_context, (size_t)nB * nT * ctx);
    alloc(ws.d_enc, (size_t)erows * H);
    alloc(ws.d_out, (size_t)nB * nT * A);
    alloc(ws.d_mask, (size_t)nB * nT);
    alloc(ws.d_enc_mask, (size_t)erows);
    alloc(ws.d_pmask, (size_t)rows);
    alloc(ws.d_t, (size_t)2 * nB);
    alloc(ws.x0, (size_t)rows * P * c.in_channels);
    alloc(ws.xn, (size_t)rows * H);
    alloc(ws.qkv, (size_t)rows * (nq + 2 * nkv));
    alloc(ws.att, (size_t)rows * nq);
    alloc(ws.act, (size_t)rows * I);
    alloc(ws.enc16, (size_t)erows * H);
    alloc(ws.encp, (size_t)erows * H);
    alloc(ws.cross_kv, (size_t)layers * erows * 2 * nkv);
    alloc(ws.x, (size_t)rows * H);
    alloc(ws.tfreq, (size_t)2 * nB * 256);
    alloc(ws.th1, (size_t)nB * H);
    alloc(ws.th2, (size_t)nB * H);
    alloc(ws.temb, (size_t)nB * H);
    alloc(ws.temb_r, (size_t)nB * H);
    alloc(ws.tproj, (size_t)nB * 6 * H);
    alloc(ws.mod, (size_t)nB * layers * 6 * H);
    alloc(ws.out_mod, (size_t)nB * 2 * H);
    alloc(ws.rope, (size_t)S * (c.head_dim / 2));
    if (!ok) {
        ws.release();
        return false;
    }
    ws.capB = nB;
    ws.capT = nT;
    ws.capE = nE;
    ws.rope_len = 0;
    return true;
}

namespace {

struct Fwd {
    const DitModel& m;
    DitWorkspace& ws;
    cudaStream_t st;
    cudaError_t rc = cudaSuccess;
    int sms;

    // D = A[rows, K] . W^T with the given epilogue; rows are treated as one sample of B*rows_per_b tokens,
    // per-sample vectors (AdaLN gate) use B / rows_per_b.
    void gemm(const __half* A, long lda, int B, int rows_per_b, const PackedWeight& W, GemmEpilogue epi, bool allow_split) {
        if (rc != cudaSuccess) return;
        GemmProblem g;
        g.A = A;
        g.B = B;
        g.L = rows_per_b;
        g.C = W.K;
        g.a_lstride = lda;
        g.a_bstride = (long)rows_per_b * lda;
        g.l_out = rows_per_b;
        g.W = &W;
        g.epi = epi;
        if (g.epi.out_bstride == 0) g.epi.out_bstride = (long)rows_per_b * g.epi.ld_out;
        g.split_k = allow_split ? 0 : 1;  // 0 = let the launcher split K when one wave of tiles would not fill the SMs
        rc = gemm_launch(g, st);
    }
};

// both embedders (t, then t - t_r) on `rows` rows of sinusoidal features laid out [2, rows, 256]: temb / temb_r [rows, H], tproj = sum
void time_embed_rows(Fwd& f, const __half* freq, __half* h1, __half* h2, float* temb, float* temb_r, float* tproj, int rows) {
    const DitModel& m = f.m;
    const int H = m.cfg.hidden_size;
    for (int which = 0; which < 2; ++which) {
        const DitTimeEmbed& te = which == 0 ? m.te : m.te_r;
        GemmEpilogue e1;
        e1.bias = te.b1;
        e1.out_f16 = h1;
        e1.silu_f16 = 1;
        e1.ld_out = H;
        f.gemm(freq + (size_t)which * rows * 256, 256, 1, rows, te.w1, e1, false);
        GemmEpilogue e2;
        e2.bias = te.b2;
        e2.out_f32 = which == 0 ? temb : temb_r;
        e2.out_f16 = h2;
        e2.silu_f16 = 1;
        e2.ld_out = H;
        f.gemm(h1, H, 1, rows, te.w2, e2, false);
        GemmEpilogue e3;
        e3.bias = te.b_proj;
        e3.out_f32 = tproj;
        e3.resid = which == 0 ? nullptr : tproj;  // proj = proj_t + proj_r
        e3.ld_out = 6 * H;
        f.gemm(h2, H, 1, rows, te.w_proj, e3, false);
    }
}

}  // namespace

cudaError_t dit_time_embed_schedule(const DitModel& m, DitWorkspace& ws, const float* t_host, int n_steps, cudaStream_t st) {
    const int H = m.cfg.hidden_size;
    if (n_steps > ws.sched_cap) {
        // grown separately from the (B, T, E) workspace: freed with it
        auto alloc = [&](auto*& ptr, size_t count) {
            void* p = nullptr;
            if (cudaMalloc(&p, count * sizeof(*ptr) + 64) != cudaSuccess) return false;
            ws.allocs.push_back(p);
            ptr = reinterpret_cast<decltype(ptr)>(p);
            return true;
        };
        const size_t n = (size_t)n_steps;
        if (!alloc(ws.sched_temb, n * H) || !alloc(ws.sched_temb_r, n * H) || !alloc(ws.sched_tproj, n * 6 * H) || !alloc(ws.sched_freq, 2 * n * 256) ||
            !alloc(ws.sched_h1, n * H) || !alloc(ws.sched_h2, n * H) || !alloc(ws.sched_t, 2 * n)) {
            ws.sched_cap = 0;
            return cudaErrorMemoryAllocation;
        }
        ws.sched_cap = n_steps;
    }
    std::vector<float> tt((size_t)2 * n_steps, 0.f);  // [t_0 .. t_{n-1} | 0 .. 0]: timestep_r == timestep in every caller
    for (int i = 0; i < n_steps; ++i) tt[(size_t)i] = t_host[i];
    cudaError_t e = cudaMemcpyAsync(ws.sched_t, tt.data(), tt.size() * 4, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return e;
    e = cudaStreamSynchronize(st);  // `tt` is pageable and goes out of scope
    if (e != cudaSuccess) return e;
    launch_timestep_features(ws.sched_t, ws.sched_freq, 2 * n_steps, st);
    Fwd f{m, ws, st, cudaSuccess, device_sm_count()};
    time_embed_rows(f, ws.sched_freq, ws.sched_h1, ws.sched_h2, ws.sched_temb, ws.sched_temb_r, ws.sched_tproj, n_steps);
    return f.rc != cudaSuccess ? f.rc : cudaGetLastError();
}

cudaError_t dit_forward_device(const DitModel& m, DitWorkspace& ws, int B, int T, int E, float* out,
                               const DitForwardOpts& opts, cudaStream_t st) {
    const DitConfig& c = m.cfg;
    const int P = c.patch_size, H = c.hidden_size, I = c.intermediate_size, A = c.audio_acoustic_hidden_dim;
    const int ctx = c.in_channels - A, hd = c.head_dim;
    const int S = (T + P - 1) / P;
    const int NQ = c.num_attention_heads, NKV = c.num_key_value_heads;
    const int nq = NQ * hd, nkv = NKV * hd;
    const int qkv_w = nq + 2 * nkv;
    const int layers_all = c.num_hidden_layers;
    const int n_layers = opts.max_layers > 0 ? std::min(opts.max_layers, layers_all) : layers_all;
    const float eps = c.rms_norm_eps;
    Fwd f{m, ws, st, cudaSuccess, device_sm_count()};

    if (ws.rope_len < S) {
        launch_build_rope_table(ws.rope, S, hd / 2, c.rope_theta, st);
        ws.rope_len = S;
    }

    // ---- inputs: pack + patch-embed
    launch_pack_dit_input(ws.d_hidden, ws.d_context, ws.x0, B, T, S, P, ctx, A, st);
    launch_patch_mask(opts.has_mask ? ws.d_mask : nullptr, ws.d_pmask, B, T, S, P, st);
    // ---- timestep embeddings -> AdaLN tables   (M = B rows; HBM-bound GEMVs through the same kernel)
    if (opts.sched_step >= 0 && opts.sched_step < ws.sched_cap) {
        const size_t r = (size_t)opts.sched_step;
        launch_build_modulation(m.layer_tables, ws.sched_tproj + r * 6 * H, ws.mod, m.out_table, ws.sched_temb + r * H, ws.sched_temb_r + r * H,
                                ws.out_mod, B, layers_all, H, st, /*shared_t=*/true);
    } else {
        launch_timestep_features(ws.d_t, ws.tfreq, 2 * B, st);
        time_embed_rows(f, ws.tfreq, ws.th1, ws.th2, ws.temb, ws.temb_r, ws.tproj, B);
        launch_build_modulation(m.layer_tables, ws.tproj, ws.mod, m.out_table, ws.temb, ws.temb_r, ws.out_mod, B, layers_all, H, st);
    }
    const long mod_bs = (long)layers_all * 6 * H;

    // ---- patch-embed and condition embedder.  They run AFTER the modulation tables on purpose: rmsnorm_mod reads its scale /
    // shift before its PDL wait, which is only safe when the kernel that wrote them is not its immediate predecessor.
    {
        GemmEpilogue e;
        e.bias = m.proj_in_b;
        e.out_f32 = ws.x;
        e.ld_out = H;
        f.gemm(ws.x0, (long)P * c.in_channels, 1, B * S, m.proj_in, e, false);
    }
    // ---- condition embedder (step-invariant: skipped when the sampler says the cache is valid)
    const bool do_cross = E > 0;
    if (do_cross && !(opts.reuse_cross_kv && ws.cross_kv_valid)) {
        launch_f32_to_f16(ws.d_enc, ws.enc16, (long)B * E * H, st);
        GemmEpilogue e;
        e.bias = m.condition_b;
        e.out_f16 = ws.encp;
        e.ld_out = H;
        f.gemm(ws.enc16, H, 1, B * E, m.condition, e, false);
        for (int i = 0; i < n_layers; ++i) {
            const DitLayer& ly = m.layers[(size_t)i];
            __half* kv = ws.cross_kv + (size_t)i * B * E * 2 * nkv;
            GemmEpilogue ek;  // K heads normed in the epilogue (no RoPE on the cross path), V stored as is
            ek.mode = EPI_HEADNORM;
            ek.hn_wa = ly.cross_kn;
            ek.hn_na = nkv;
            ek.hn_eps = eps;
            ek.out_f16 = kv;
            ek.ld_out = 2 * nkv;
            f.gemm(ws.encp, H, 1, B * E, ly.cross_kv, ek, false);
        }
    }

    // ---- transformer layers
    for (int i = 0; i < n_layers; ++i) {
        const DitLayer& ly = m.layers[(size_t)i];
        const float* mod = ws.mod + (size_t)i * 6 * H;  // + b * mod_bs
        const float* shift_msa = mod + 0 * H;
        const float* scale_msa = mod + 1 * H;
        const float* gate_msa = mod + 2 * H;
        const float* c_shift = mod + 3 * H;
        const float* c_scale = mod + 4 * H;
        const float* c_gate = mod + 5 * H;

        // self-attention
        // early_mod: the tables were written by build_modulation ahead of the patch-embed GEMM, and every kernel since waits before
        // it triggers its dependents, so no rmsnorm can start before that writer has completed
        launch_rmsnorm_mod(ws.x, ly.self_norm, scale_msa, shift_msa, mod_bs, ws.xn, B, S, H, eps, st, nullptr, true);
        {
            GemmEpilogue e;  // fused: per-head RMSNorm(q, k) + NEOX RoPE at position l; V stored as is
            e.mode = EPI_HEADNORM;
            e.hn_wa = ly.self_qn;
            e.hn_na = nq;
            e.hn_wb = ly.self_kn;
            e.hn_nb = nkv;
            e.rope = ws.rope;
            e.hn_eps = eps;
            e.out_f16 = ws.qkv;
            e.ld_out = qkv_w;
            f.gemm(ws.xn, H, B, S, ly.self_qkv, e, false);
        }
        launch_attention(ws.qkv, qkv_w, ws.qkv + nq, qkv_w, ws.qkv + nq + nkv, qkv_w, ws.att, nq, ws.d_pmask, B, S, S, NQ,
                         NKV, ly.sliding ? c.sliding_window : -1, st);
        {
            GemmEpilogue e;
            e.gate = gate_msa;
            e.gate_bstride = mod_bs;
            e.resid = ws.x;
            e.out_f32 = ws.x;
            e.ld_out = H;
            f.gemm(ws.att, nq, B, S, ly.self_o, e, true);
        }
        // cross-attention
        if (do_cross) {
            launch_rmsnorm_mod(ws.x, ly.cross_norm, nullptr, nullptr, 0, ws.xn, B, S, H, eps, st);
            {
                GemmEpilogue e;
                e.mode = EPI_HEADNORM;
                e.hn_wa = ly.cross_qn;
                e.hn_na = nq;
                e.hn_eps = eps;
                e.out_f16 = ws.qkv;
                e.ld_out = nq;
                f.gemm(ws.xn, H, 1, B * S, ly.cross_q, e, false);
            }
            const __half* kv = ws.cross_kv + (size_t)i * B * E * 2 * nkv;
            launch_attention(ws.qkv, nq, kv, 2 * nkv, kv + nkv, 2 * nkv, ws.att, nq, opts.has_enc_mask ? ws.d_enc_mask : nullptr, B,
                             S, E, NQ, NKV, -1, st);
            {
                GemmEpilogue e;
                e.resid = ws.x;
                e.out_f32 = ws.x;
                e.ld_out = H;
                f.gemm(ws.att, nq, B, S, ly.cross_o, e, true);
            }
        }
        // MLP
        launch_rmsnorm_mod(ws.x, ly.mlp_norm, c_scale, c_shift, mod_bs, ws.xn, B, S, H, eps, st, nullptr, true);
        {
            GemmEpilogue e;
            e.mode = EPI_SWIGLU;
            e.out_f16 = ws.act;
            e.ld_out = I;
            f.gemm(ws.xn, H, 1, B * S, ly.mlp_gate_up, e, false);
        }
        {
            GemmEpilogue e;
            e.gate = c_gate;
            e.gate_bstride = mod_bs;
            e.resid = ws.x;
            e.out_f32 = ws.x;
            e.ld_out = H;
            f.gemm(ws.act, I, B, S, ly.mlp_down, e, true);
        }
    }

    // ---- output head: AdaLN(norm_out) -> proj_out (un-patch) -> crop to T
    launch_rmsnorm_mod(ws.x, m.norm_out, ws.out_mod + H, ws.out_mod, 2L * H, ws.xn, B, S, H, eps, st, nullptr, n_layers > 0);
    {
        GemmEpilogue e;
        e.bias = m.proj_out_b;
        e.out_f32 = out;
        e.ld_out = (long)P * A;
        e.out_bstride = (long)T * A;
        e.out_valid = (long)T * A;
        f.gemm(ws.xn, H, B, S, m.proj_out, e, false);
    }
    if (f.rc != cudaSuccess) return f.rc;
    ws.cross_kv_valid = do_cross;
    return cudaGetLastError();
}

}  // namespace ace
