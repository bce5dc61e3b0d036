// Fused attention for the DiT (see kernels.h): one pass, no [k_len, q_len, heads] score tensor, no host mask.
//   full / sliding-window (banded: only the key blocks inside the band are visited) / cross, GQA by index,
//   key-padding mask, head_dim 128, fp16 operands, fp32 softmax state and accumulators.
// CTA = 4 warps x 16 query rows; K/V tiles of 64 keys double-buffered in shared memory with cp.async,
// XOR-swizzled rows so ldmatrix is bank-conflict free; tensor-core m16n8k16 MMAs.
// reference semantics: acestep_dit_model.cpp:1132-1173 (mask rule), 1212-1256 (scale, softmax, PV, GQA repeat)
#include "kernels.h"

#include <cstdint>
#include <cstdio>

namespace ace {

namespace {

constexpr int HD = 128;       // head_dim
constexpr int BQ = 64;        // query rows per CTA
constexpr int BKEY = 64;      // keys per tile
constexpr int ATT_THREADS = 128;
constexpr int TILE_BYTES = BKEY * HD * 2;  // 16 KB
constexpr int ATT_SMEM = TILE_BYTES * 5 + 2 * BKEY * 4;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint32_t swz(int row, int chunk) { return (uint32_t)(row * 256 + ((chunk ^ (row & 7)) << 4)); }

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
    const int sz = valid ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}

// load a [64 x 128] fp16 tile (rows r0.. of a token-major tensor) into swizzled smem; rows >= n_rows are zero
__device__ __forceinline__ void load_tile(uint32_t sdst, const __half* base, long stride, int r0, int n_rows) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int idx = threadIdx.x + ATT_THREADS * i;
        const int row = idx >> 4, chunk = idx & 15;
        const bool ok = (r0 + row) < n_rows;
        const __half* src = base + (long)(ok ? (r0 + row) : 0) * stride + chunk * 8;
        cp_async16(sdst + swz(row, chunk), src, ok);
    }
}

__global__ void __launch_bounds__(ATT_THREADS)
attention_kernel(const __half* __restrict__ Q, long q_stride, const __half* __restrict__ K, long k_stride,
                 const __half* __restrict__ V, long v_stride, __half* __restrict__ O, long o_stride,
                 const int32_t* __restrict__ key_mask, int q_len, int k_len, int n_rep, int window, int causal) {
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t sQ = smem_u32(smem);
    const uint32_t sK = sQ + TILE_BYTES;       // 2 stages
    const uint32_t sV = sK + 2 * TILE_BYTES;   // 2 stages
    float* kbias = reinterpret_cast<float*>(smem + 5 * TILE_BYTES);  // [2][64] 0 or -inf

    const int q0 = blockIdx.x * BQ;
    const int h = blockIdx.y;
    const int b = blockIdx.z;
    const int hk = h / n_rep;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, t4 = lane & 3;

    const __half* Qb = Q + (long)b * q_len * q_stride + h * HD;
    const __half* Kb = K + (long)b * k_len * k_stride + hk * HD;
    const __half* Vb = V + (long)b * k_len * v_stride + hk * HD;
    const int32_t* mb = key_mask ? key_mask + (long)b * k_len : nullptr;

    int kb_lo = 0, kb_hi = (k_len + BKEY - 1) / BKEY - 1;
    if (window >= 0) {
        kb_lo = max(0, (q0 - window) / BKEY);
        kb_hi = min(kb_hi, (min(q0 + BQ - 1, q_len - 1) + window) / BKEY);
    }
    if (causal) kb_hi = min(kb_hi, min(q0 + BQ - 1, q_len - 1) / BKEY);  // keys after the last query of the tile are never visible

    auto issue_kv = [&](int kb, int stage) {
        load_tile(sK + stage * TILE_BYTES, Kb, k_stride, kb * BKEY, k_len);
        load_tile(sV + stage * TILE_BYTES, Vb, v_stride, kb * BKEY, k_len);
        if (threadIdx.x < BKEY) {
            const int k = kb * BKEY + threadIdx.x;
            const bool ok = k < k_len && (!mb || mb[k] != 0);
            kbias[stage * BKEY + threadIdx.x] = ok ? 0.f : -INFINITY;
        }
    };

    load_tile(sQ, Qb, q_stride, q0, q_len);
    issue_kv(kb_lo, 0);
    cp_async_commit();
    cp_async_wait<0>();
    __syncthreads();

    // Q fragments: 8 k-steps of 16 dims
    uint32_t qf[8][4];
    {
        const int row = warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) ldsm_x4(sQ + swz(row, ks * 2 + (lane >> 4)), qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3]);
    }

    float o[16][4];
#pragma unroll
    for (int i = 0; i < 16; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
    float m_run[2] = {-INFINITY, -INFINITY};
    float l_run[2] = {0.f, 0.f};
    const float sc = 0.08838834764831845f * 1.4426950408889634f;  // 1/sqrt(128) * log2(e)
    const int qrow[2] = {q0 + warp * 16 + g, q0 + warp * 16 + g + 8};

    for (int kb = kb_lo; kb <= kb_hi; ++kb) {
        const int stage = (kb - kb_lo) & 1;
        if (kb + 1 <= kb_hi) {  // prefetch next tile into the other stage (its previous reader finished: trailing barrier)
            issue_kv(kb + 1, stage ^ 1);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();

        // ---- S = Q K^T  (16 x 64 per warp)
        float s[8][4];
#pragma unroll
        for (int i = 0; i < 8; ++i) { s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f; }
        const uint32_t kbase = sK + stage * TILE_BYTES;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
#pragma unroll
            for (int np = 0; np < 4; ++np) {  // pairs of 8-key tiles
                uint32_t b0, b1, b2, b3;
                const int key = np * 16 + (lane & 7) + (lane >> 4) * 8;
                ldsm_x4(kbase + swz(key, ks * 2 + ((lane >> 3) & 1)), b0, b1, b2, b3);
                mma16816(s[2 * np], qf[ks], b0, b1);
                mma16816(s[2 * np + 1], qf[ks], b2, b3);
            }
        }

        // ---- mask + online softmax (rows g and g+8 of this warp's 16)
        const float* kbz = kbias + stage * BKEY;
        float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int kl = nt * 8 + 2 * t4 + (e & 1);
                const int r = e >> 1;
                float v = s[nt][e] * sc + kbz[kl];
                if (window >= 0 || causal) {
                    const int d = qrow[r] - (kb * BKEY + kl);
                    if (window >= 0 && (d > window || d < -window)) v = -INFINITY;
                    if (causal && d < 0) v = -INFINITY;  // qwen_model.cpp:611-613 / acestep_dit_model.cpp:1148-1150
                }
                s[nt][e] = v;
                mx[r] = fmaxf(mx[r], v);
            }
        }
        float alpha[2], msafe[2];
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
            mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
            const float mnew = fmaxf(m_run[r], mx[r]);
            msafe[r] = (mnew == -INFINITY) ? 0.f : mnew;
            alpha[r] = exp2f(m_run[r] - msafe[r]);  // m_run = -inf -> 0
            m_run[r] = mnew;
            l_run[r] *= alpha[r];
        }
        uint32_t pf[4][4];  // P as A fragments: 4 k-steps of 16 keys
        float psum[2] = {0.f, 0.f};
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            const float p0 = exp2f(s[nt][0] - msafe[0]);
            const float p1 = exp2f(s[nt][1] - msafe[0]);
            const float p2 = exp2f(s[nt][2] - msafe[1]);
            const float p3 = exp2f(s[nt][3] - msafe[1]);
            psum[0] += p0 + p1;
            psum[1] += p2 + p3;
            const int kk = nt >> 1;
            if ((nt & 1) == 0) {
                pf[kk][0] = pack_h2(p0, p1);
                pf[kk][1] = pack_h2(p2, p3);
            } else {
                pf[kk][2] = pack_h2(p0, p1);
                pf[kk][3] = pack_h2(p2, p3);
            }
        }
        l_run[0] += psum[0];
        l_run[1] += psum[1];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            o[i][0] *= alpha[0]; o[i][1] *= alpha[0];
            o[i][2] *= alpha[1]; o[i][3] *= alpha[1];
        }

        // ---- O += P V
        const uint32_t vbase = sV + stage * TILE_BYTES;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
            for (int dp = 0; dp < 8; ++dp) {  // pairs of 8-dim tiles
                uint32_t b0, b1, b2, b3;
                const int key = kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
                ldsm_x4_t(vbase + swz(key, dp * 2 + (lane >> 4)), b0, b1, b2, b3);
                mma16816(o[2 * dp], pf[kk], b0, b1);
                mma16816(o[2 * dp + 1], pf[kk], b2, b3);
            }
        }
        __syncthreads();  // everyone done with this stage before it is refilled
    }

    // ---- finalise: O / l, write token-major fp16
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        float l = l_run[r];
        l += __shfl_xor_sync(0xffffffffu, l, 1);
        l += __shfl_xor_sync(0xffffffffu, l, 2);
        const float inv = l > 0.f ? 1.0f / l : 0.f;
        const int q = qrow[r];
        if (q < q_len) {
            __half* op = O + ((long)b * q_len + q) * o_stride + h * HD;
#pragma unroll
            for (int dt = 0; dt < 16; ++dt) {
                const uint32_t v = pack_h2(o[dt][2 * r] * inv, o[dt][2 * r + 1] * inv);
                *reinterpret_cast<uint32_t*>(op + dt * 8 + 2 * t4) = v;
            }
        }
    }
}

}  // namespace

void launch_attention(const __half* Q, long q_stride, const __half* K, long k_stride, const __half* V, long v_stride,
                      __half* O, long o_stride, const int32_t* key_mask, int B, int q_len, int k_len, int n_heads,
                      int n_kv_heads, int window, cudaStream_t st, bool causal) {
    note_launch();
    const double k_eff = window >= 0 ? (double)(2 * window + 1 < k_len ? 2 * window + 1 : k_len) : (double)k_len;
    ProfScope prof_scope(PROF_ATTN, 4.0 * B * n_heads * (double)q_len * k_eff * 128.0, 0.0, st);
    if (attention_use_tc() && !causal) {  // the causal mask (text encoder, once per song) lives in the mma.sync kernel only
        const cudaError_t e = launch_attention_tc(Q, q_stride, K, k_stride, V, v_stride, O, o_stride, key_mask, B, q_len, k_len, n_heads,
                                                  n_kv_heads, window, st);
        if (e == cudaSuccess) return;
        std::fprintf(stderr, "[ace_b200] tcgen05 attention launch failed (%s); using the mma.sync kernel\n", cudaGetErrorString(e));
    }
    static DeviceOnce attr;
    attr.run([](int&) { return cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_SMEM); });
    dim3 grid((q_len + BQ - 1) / BQ, n_heads, B);
    attention_kernel<<<grid, ATT_THREADS, ATT_SMEM, st>>>(Q, q_stride, K, k_stride, V, v_stride, O, o_stride, key_mask,
                                                          q_len, k_len, n_heads / n_kv_heads, window, causal ? 1 : 0);
}

}  // namespace ace
