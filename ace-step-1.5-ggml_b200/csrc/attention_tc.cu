// Fused attention on the 5th-gen tensor cores (tcgen05 + TMEM + TMA), head_dim 128, fp16 operands, fp32 softmax.
// Same contract as attention.cu (kernels.h: launch_attention): full / sliding-window (only key tiles inside the band are
// visited) / cross, GQA by head index, key-padding mask.   reference semantics: acestep_dit_model.cpp:1132-1173, 1212-1256
//
// CTA = 128 queries of one (sample, head), 6 warps, TWO CTAs per SM (112 KB smem, 256 TMEM columns each) so one CTA's
// softmax overlaps the other's MMAs:
//   warp 4    TMA producer: Q once; K / V tiles of 64 keys, 2 stages; key-padding bits of the tile (one 64-bit word)
//   warp 5    TMEM allocator + single-thread MMA issuer
//               S_j  = Q . K_j^T      M=128 N=64  K=128   -> TMEM S[j&1]      (double-buffered: S_{j+1} runs during softmax_j)
//               O   += P_j . V_j      M=128 N=128 K=64    -> TMEM O           (V is the MN-major B operand, as TMA lands it)
//   warps 0-3 softmax: thread = one query row = one TMEM lane (no shuffles): tcgen05.ld S, scale + mask, running max,
//             p = 2^(s - m), row sum, P -> fp16 -> swizzled smem (A operand of the second MMA).
// O stays in TMEM across key tiles. The reference max is moved LAZILY: only when a row's max grows by more than 2^8 is O
// (and the row sum) rescaled through tcgen05.ld/st; otherwise p is simply formed against the stale max (<= 2^8, exact in
// fp16 range) — the result O / l does not depend on which reference max was used.
#include "kernels.h"
#include "ptx.cuh"
#include "tmap.h"

#include <cstdint>
#include <cstdlib>
#include <string>

namespace ace {

namespace {

constexpr int HD = 128;
constexpr int BQ = 128;
constexpr int BKV = 64;
constexpr int Q_BYTES = BQ * HD * 2;          // 32 KB: two 64-column slabs of [128 x 128 B]
constexpr int KV_TILE = BKV * HD * 2;         // 16 KB: two slabs of [64 x 128 B]
constexpr int P_BYTES = BQ * BKV * 2;         // 16 KB
constexpr int KV_STAGES = 2;
constexpr int OFF_K = Q_BYTES;
constexpr int OFF_V = OFF_K + KV_STAGES * KV_TILE;
constexpr int OFF_P = OFF_V + KV_STAGES * KV_TILE;
constexpr int OFF_BAR = OFF_P + P_BYTES;      // 114 688
constexpr int ATT_TC_SMEM = OFF_BAR + 256;
constexpr int ATT_TC_THREADS = 192;
constexpr int TMEM_COLS = 256;                // S0 [0,64) S1 [64,128) O [128,256)
constexpr float RESCALE_THRESHOLD = 8.0f;     // log2 units

struct AttnParams {
    __half* O;
    long o_stride;
    const int32_t* key_mask;
    int q_len, k_len, n_rep, window;
};

__global__ void __launch_bounds__(ATT_TC_THREADS, 2)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                    const __grid_constant__ CUtensorMap tmV, const AttnParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sQ = smem;
    uint8_t* sK = smem + OFF_K;
    uint8_t* sV = smem + OFF_V;
    uint8_t* sP = smem + OFF_P;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
    uint64_t* q_full = bars;            // [1]
    uint64_t* kv_full = bars + 1;       // [2]
    uint64_t* kv_empty = bars + 3;      // [2]
    uint64_t* s_ready = bars + 5;       // [2]
    uint64_t* s_free = bars + 7;        // [2]
    uint64_t* p_ready = bars + 9;       // [1]
    uint64_t* pv_done = bars + 10;      // [1]
    uint64_t* kmask = bars + 12;        // [2] bit i = key (tile start + i) is visible
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 14);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q0 = blockIdx.x * BQ;
    const int h = blockIdx.y, b = blockIdx.z;
    const int hk = h / p.n_rep;

    int kb_lo = 0, kb_hi = (p.k_len + BKV - 1) / BKV - 1;
    if (p.window >= 0) {
        kb_lo = max(0, (q0 - p.window) / BKV);
        kb_hi = min(kb_hi, (min(q0 + BQ - 1, p.q_len - 1) + p.window) / BKV);
    }
    const int n_tiles = kb_hi - kb_lo + 1;

    if (threadIdx.x == 0) {
        if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();  // the 128-byte swizzle needs 1 KB aligned tiles
        ptx::mbar_init(q_full, 1);
        for (int s = 0; s < 2; ++s) {
            ptx::mbar_init(&kv_full[s], 1);
            ptx::mbar_init(&kv_empty[s], 1);
            ptx::mbar_init(&s_ready[s], 1);
            ptx::mbar_init(&s_free[s], 4);
        }
        ptx::mbar_init(p_ready, 4);
        ptx::mbar_init(pv_done, 1);
        ptx::fence_mbar_init();
    }
    if (warp == 5) ptx::tmem_alloc<TMEM_COLS>(tmem_slot);
    if (warp == 4 && lane == 0) {
        ptx::prefetch_tmap(&tmQ);
        ptx::prefetch_tmap(&tmK);
        ptx::prefetch_tmap(&tmV);
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    ptx::pdl_wait();
    ptx::pdl_launch();

    if (warp == 4) {
        // ===================== producer =====================
        if (lane == 0) {
            ptx::mbar_expect_tx(q_full, Q_BYTES);
            ptx::tma_load_4d(sQ, &tmQ, q_full, 0, h, q0, b);
            ptx::tma_load_4d(sQ + Q_BYTES / 2, &tmQ, q_full, 64, h, q0, b);
        }
        const int32_t* mb = p.key_mask ? p.key_mask + (long)b * p.k_len : nullptr;
        for (int j = 0; j < n_tiles; ++j) {
            const int s = j & 1;
            const uint32_t ph = (j >> 1) & 1;
            const int k0 = (kb_lo + j) * BKV;
            ptx::mbar_wait(&kv_empty[s], ph ^ 1);
            const int ka = k0 + lane, kb2 = k0 + 32 + lane;
            const bool ok_a = ka < p.k_len && (!mb || mb[ka] != 0);
            const bool ok_b = kb2 < p.k_len && (!mb || mb[kb2] != 0);
            const uint32_t lo = __ballot_sync(0xffffffffu, ok_a), hi = __ballot_sync(0xffffffffu, ok_b);
            if (lane == 0) {
                kmask[s] = (uint64_t)lo | ((uint64_t)hi << 32);
                ptx::mbar_expect_tx(&kv_full[s], 2 * KV_TILE);
                ptx::tma_load_4d(sK + s * KV_TILE, &tmK, &kv_full[s], 0, hk, k0, b);
                ptx::tma_load_4d(sK + s * KV_TILE + KV_TILE / 2, &tmK, &kv_full[s], 64, hk, k0, b);
                ptx::tma_load_4d(sV + s * KV_TILE, &tmV, &kv_full[s], 0, hk, k0, b);
                ptx::tma_load_4d(sV + s * KV_TILE + KV_TILE / 2, &tmV, &kv_full[s], 64, hk, k0, b);
            }
            __syncwarp();
        }
    } else if (warp == 5) {
        // ===================== MMA issuer =====================
        constexpr uint32_t idesc_s = ptx::umma_idesc_f16(BQ, BKV);
        constexpr uint32_t idesc_o = ptx::umma_idesc_f16_bmn(BQ, HD);
        const uint32_t q_addr = ptx::smem_u32(sQ), p_addr = ptx::smem_u32(sP);
        auto issue_s = [&](int j) {  // S[j&1] = Q . K_j^T
            const int s = j & 1;
            const uint32_t k_addr = ptx::smem_u32(sK + s * KV_TILE);
#pragma unroll
            for (int kk = 0; kk < HD / 16; ++kk) {
                const uint64_t da = ptx::umma_desc_sw128(q_addr + (kk >> 2) * (Q_BYTES / 2) + (kk & 3) * 32);
                const uint64_t db = ptx::umma_desc_sw128(k_addr + (kk >> 2) * (KV_TILE / 2) + (kk & 3) * 32);
                ptx::umma_f16(tmem_base + (uint32_t)(s * BKV), da, db, idesc_s, kk > 0 ? 1u : 0u);
            }
            ptx::umma_commit(&s_ready[s]);
        };
        ptx::mbar_wait(q_full, 0);
        ptx::mbar_wait(&kv_full[0], 0);
        ptx::tc_fence_after();
        if (lane == 0) issue_s(0);
        __syncwarp();
        for (int j = 0; j < n_tiles; ++j) {
            if (j + 1 < n_tiles) {  // next tile's scores while the softmax warps work on tile j
                const int s1 = (j + 1) & 1;
                const uint32_t u = (uint32_t)((j + 1) >> 1);
                ptx::mbar_wait(&kv_full[s1], u & 1);
                ptx::mbar_wait(&s_free[s1], (u & 1) ^ 1);
                ptx::tc_fence_after();
                if (lane == 0) issue_s(j + 1);
                __syncwarp();
            }
            ptx::mbar_wait(p_ready, j & 1);
            ptx::tc_fence_after();
            if (lane == 0) {
                const uint32_t v_addr = ptx::smem_u32(sV + (j & 1) * KV_TILE);
#pragma unroll
                for (int kk = 0; kk < BKV / 16; ++kk) {
                    const uint64_t da = ptx::umma_desc_sw128(p_addr + kk * 32);
                    const uint64_t db = ptx::umma_desc_mn_sw128(v_addr + kk * 2048, KV_TILE / 2, 1024);
                    ptx::umma_f16(tmem_base + 128u, da, db, idesc_o, (j > 0 || kk > 0) ? 1u : 0u);
                }
                ptx::umma_commit(&kv_empty[j & 1]);
                ptx::umma_commit(pv_done);
            }
            __syncwarp();
        }
    } else {
        // ===================== softmax (warps 0..3): thread = query row = TMEM lane =====================
        const int r = warp * 32 + lane;
        const int q = q0 + r;
        const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
        const float sc = 0.08838834764831845f * 1.4426950408889634f;  // 1/sqrt(128) * log2(e)
        float m_used = -INFINITY, l = 0.f;
        for (int j = 0; j < n_tiles; ++j) {
            const int s = j & 1;
            const int k0 = (kb_lo + j) * BKV;
            ptx::mbar_wait(&s_ready[s], (j >> 1) & 1);
            ptx::tc_fence_after();
            uint32_t r0[32], r1[32];
            ptx::tmem_ld32(lane_addr + (uint32_t)(s * BKV), r0);
            ptx::tmem_ld32(lane_addr + (uint32_t)(s * BKV + 32), r1);
            ptx::tmem_ld_wait();
            const uint64_t km = kmask[s];
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&s_free[s]);

            float v[64];
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                v[i] = __uint_as_float(r0[i]) * sc;
                v[32 + i] = __uint_as_float(r1[i]) * sc;
            }
            const bool band_edge = p.window >= 0 && ((q0 + BQ - 1) - k0 > p.window || (k0 + BKV - 1) - q0 > p.window);
            if (km != ~0ull) {  // padded / out-of-range keys (warp-uniform)
                const uint32_t klo = (uint32_t)km, khi = (uint32_t)(km >> 32);
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    if (!((klo >> i) & 1u)) v[i] = -INFINITY;
                    if (!((khi >> i) & 1u)) v[32 + i] = -INFINITY;
                }
            }
            if (band_edge) {
#pragma unroll
                for (int i = 0; i < 64; ++i) {
                    const int d = q - (k0 + i);
                    if (d > p.window || d < -p.window) v[i] = -INFINITY;
                }
            }
            float mx = v[0];
#pragma unroll
            for (int i = 1; i < 64; ++i) mx = fmaxf(mx, v[i]);
            const float m_new = fmaxf(m_used, mx);
            const bool need = m_new > m_used + RESCALE_THRESHOLD;  // also true when m_used = -inf and the row saw a key
            const float m_next = need ? m_new : m_used;
            const float alpha = need ? ptx::ex2(m_used - m_new) : 1.0f;  // m_used = -inf -> 0 (O row and l are 0 anyway)
            const float msafe = (m_next == -INFINITY) ? 0.f : m_next;
            uint32_t ph[32];
            float psum = 0.f;
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const float p0 = ptx::ex2(v[2 * i] - msafe), p1 = ptx::ex2(v[2 * i + 1] - msafe);
                psum += p0 + p1;
                const __half2 hh = __floats2half2_rn(p0, p1);
                ph[i] = *reinterpret_cast<const uint32_t*>(&hh);
            }
            // P buffer free and O quiescent once the previous tile's second MMA has completed
            if (j > 0) {
                ptx::mbar_wait(pv_done, (j - 1) & 1);
                ptx::tc_fence_after();
                if (__any_sync(0xffffffffu, need)) {
#pragma unroll 1
                    for (int c = 0; c < 4; ++c) {
                        uint32_t o[32];
                        ptx::tmem_ld32(lane_addr + 128u + (uint32_t)(c * 32), o);
                        ptx::tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                        ptx::tmem_st32(lane_addr + 128u + (uint32_t)(c * 32), o);
                    }
                    ptx::tmem_st_wait();
                }
            }
            l = l * alpha + psum;
            m_used = m_next;
#pragma unroll
            for (int c = 0; c < 8; ++c)
                *reinterpret_cast<uint4*>(sP + r * 128 + ((c ^ (r & 7)) << 4)) = make_uint4(ph[4 * c], ph[4 * c + 1], ph[4 * c + 2], ph[4 * c + 3]);
            ptx::fence_proxy_async_smem();
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(p_ready);
        }
        // ---- finalise: O / l -> fp16, token-major
        ptx::mbar_wait(pv_done, (n_tiles - 1) & 1);
        ptx::tc_fence_after();
        const float inv = l > 0.f ? 1.0f / l : 0.f;
        __half* op = p.O + ((long)b * p.q_len + q) * p.o_stride + (long)h * HD;
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
            uint32_t o[32];
            ptx::tmem_ld32(lane_addr + 128u + (uint32_t)(c * 32), o);
            ptx::tmem_ld_wait();
            if (q < p.q_len) {
                uint32_t w[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const __half2 hh = __floats2half2_rn(__uint_as_float(o[2 * i]) * inv, __uint_as_float(o[2 * i + 1]) * inv);
                    w[i] = *reinterpret_cast<const uint32_t*>(&hh);
                }
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    *reinterpret_cast<uint4*>(op + c * 32 + i * 8) = make_uint4(w[4 * i], w[4 * i + 1], w[4 * i + 2], w[4 * i + 3]);
            }
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 5) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<TMEM_COLS>(tmem_base);
    }
}

// ---------------------------------------------------------------- 8 softmax warps: two threads per query row
// Same pipeline as attention_tc_kernel, but the 64 score columns of a row are split between two threads (warp w and w + 4
// share a TMEM lane quadrant; w < 4 takes keys 0..31 of the tile, w >= 4 keys 32..63), halving the serial exp / pack chain
// that bounded the 4-warp version (ncu: tensor pipe 30 %, issue 31 %, MUFU 33 % — a latency chain, nothing saturated).
// Both threads of a row must exponentiate against the SAME reference max.  It only moves in the lazy-rescale events, so the
// common path needs one OR-reduction barrier per tile ("does any row need to move its max?") and no data exchange; the rare
// path swaps the two half-row maxima through the (idle) P buffer, and each thread rescales its half of the O columns.
constexpr int ATT8_THREADS = 320;

__global__ void __launch_bounds__(ATT8_THREADS, 2)
attention_tc8_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                     const __grid_constant__ CUtensorMap tmV, const AttnParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sQ = smem;
    uint8_t* sK = smem + OFF_K;
    uint8_t* sV = smem + OFF_V;
    uint8_t* sP = smem + OFF_P;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
    uint64_t* q_full = bars;
    uint64_t* kv_full = bars + 1;
    uint64_t* kv_empty = bars + 3;
    uint64_t* s_ready = bars + 5;
    uint64_t* s_free = bars + 7;
    uint64_t* p_ready = bars + 9;
    uint64_t* pv_done = bars + 10;
    uint64_t* kmask = bars + 12;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 14);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q0 = blockIdx.x * BQ;
    const int h = blockIdx.y, b = blockIdx.z;
    const int hk = h / p.n_rep;

    int kb_lo = 0, kb_hi = (p.k_len + BKV - 1) / BKV - 1;
    if (p.window >= 0) {
        kb_lo = max(0, (q0 - p.window) / BKV);
        kb_hi = min(kb_hi, (min(q0 + BQ - 1, p.q_len - 1) + p.window) / BKV);
    }
    const int n_tiles = kb_hi - kb_lo + 1;

    if (threadIdx.x == 0) {
        if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();
        ptx::mbar_init(q_full, 1);
        for (int s = 0; s < 2; ++s) {
            ptx::mbar_init(&kv_full[s], 1);
            ptx::mbar_init(&kv_empty[s], 1);
            ptx::mbar_init(&s_ready[s], 1);
            ptx::mbar_init(&s_free[s], 8);
        }
        ptx::mbar_init(p_ready, 8);
        ptx::mbar_init(pv_done, 1);
        ptx::fence_mbar_init();
    }
    if (warp == 9) ptx::tmem_alloc<TMEM_COLS>(tmem_slot);
    if (warp == 8 && lane == 0) {
        ptx::prefetch_tmap(&tmQ);
        ptx::prefetch_tmap(&tmK);
        ptx::prefetch_tmap(&tmV);
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    ptx::pdl_wait();
    ptx::pdl_launch();

    if (warp == 8) {
        // ===================== producer =====================
        if (lane == 0) {
            ptx::mbar_expect_tx(q_full, Q_BYTES);
            ptx::tma_load_4d(sQ, &tmQ, q_full, 0, h, q0, b);
            ptx::tma_load_4d(sQ + Q_BYTES / 2, &tmQ, q_full, 64, h, q0, b);
        }
        const int32_t* mb = p.key_mask ? p.key_mask + (long)b * p.k_len : nullptr;
        for (int j = 0; j < n_tiles; ++j) {
            const int s = j & 1;
            const uint32_t ph = (j >> 1) & 1;
            const int k0 = (kb_lo + j) * BKV;
            ptx::mbar_wait(&kv_empty[s], ph ^ 1);
            const int ka = k0 + lane, kb2 = k0 + 32 + lane;
            const bool ok_a = ka < p.k_len && (!mb || mb[ka] != 0);
            const bool ok_b = kb2 < p.k_len && (!mb || mb[kb2] != 0);
            const uint32_t lo = __ballot_sync(0xffffffffu, ok_a), hi = __ballot_sync(0xffffffffu, ok_b);
            if (lane == 0) {
                kmask[s] = (uint64_t)lo | ((uint64_t)hi << 32);
                ptx::mbar_expect_tx(&kv_full[s], 2 * KV_TILE);
                ptx::tma_load_4d(sK + s * KV_TILE, &tmK, &kv_full[s], 0, hk, k0, b);
                ptx::tma_load_4d(sK + s * KV_TILE + KV_TILE / 2, &tmK, &kv_full[s], 64, hk, k0, b);
                ptx::tma_load_4d(sV + s * KV_TILE, &tmV, &kv_full[s], 0, hk, k0, b);
                ptx::tma_load_4d(sV + s * KV_TILE + KV_TILE / 2, &tmV, &kv_full[s], 64, hk, k0, b);
            }
            __syncwarp();
        }
    } else if (warp == 9) {
        // ===================== MMA issuer (identical to the 4-warp kernel) =====================
        constexpr uint32_t idesc_s = ptx::umma_idesc_f16(BQ, BKV);
        constexpr uint32_t idesc_o = ptx::umma_idesc_f16_bmn(BQ, HD);
        const uint32_t q_addr = ptx::smem_u32(sQ), p_addr = ptx::smem_u32(sP);
        auto issue_s = [&](int j) {
            const int s = j & 1;
            const uint32_t k_addr = ptx::smem_u32(sK + s * KV_TILE);
#pragma unroll
            for (int kk = 0; kk < HD / 16; ++kk) {
                const uint64_t da = ptx::umma_desc_sw128(q_addr + (kk >> 2) * (Q_BYTES / 2) + (kk & 3) * 32);
                const uint64_t db = ptx::umma_desc_sw128(k_addr + (kk >> 2) * (KV_TILE / 2) + (kk & 3) * 32);
                ptx::umma_f16(tmem_base + (uint32_t)(s * BKV), da, db, idesc_s, kk > 0 ? 1u : 0u);
            }
            ptx::umma_commit(&s_ready[s]);
        };
        ptx::mbar_wait(q_full, 0);
        ptx::mbar_wait(&kv_full[0], 0);
        ptx::tc_fence_after();
        if (lane == 0) issue_s(0);
        __syncwarp();
        for (int j = 0; j < n_tiles; ++j) {
            if (j + 1 < n_tiles) {
                const int s1 = (j + 1) & 1;
                const uint32_t u = (uint32_t)((j + 1) >> 1);
                ptx::mbar_wait(&kv_full[s1], u & 1);
                ptx::mbar_wait(&s_free[s1], (u & 1) ^ 1);
                ptx::tc_fence_after();
                if (lane == 0) issue_s(j + 1);
                __syncwarp();
            }
            ptx::mbar_wait(p_ready, j & 1);
            ptx::tc_fence_after();
            if (lane == 0) {
                const uint32_t v_addr = ptx::smem_u32(sV + (j & 1) * KV_TILE);
#pragma unroll
                for (int kk = 0; kk < BKV / 16; ++kk) {
                    const uint64_t da = ptx::umma_desc_sw128(p_addr + kk * 32);
                    const uint64_t db = ptx::umma_desc_mn_sw128(v_addr + kk * 2048, KV_TILE / 2, 1024);
                    ptx::umma_f16(tmem_base + 128u, da, db, idesc_o, (j > 0 || kk > 0) ? 1u : 0u);
                }
                ptx::umma_commit(&kv_empty[j & 1]);
                ptx::umma_commit(pv_done);
            }
            __syncwarp();
        }
    } else {
        // ===================== softmax: warps 0..7, thread = (query row, half of the tile's keys) =====================
        const int quad = warp & 3, hf = warp >> 2;
        const int r = quad * 32 + lane;
        const int q = q0 + r;
        const uint32_t lane_addr = tmem_base + ((uint32_t)(quad * 32) << 16);
        const float sc = 0.08838834764831845f * 1.4426950408889634f;
        float* xch = reinterpret_cast<float*>(sP);  // [2][128] scratch inside the idle P buffer (rare path / final sum)
        float m_used = -INFINITY, l = 0.f;
        for (int j = 0; j < n_tiles; ++j) {
            const int s = j & 1;
            const int k0 = (kb_lo + j) * BKV + hf * 32;   // first key of this thread's 32 columns
            ptx::mbar_wait(&s_ready[s], (j >> 1) & 1);
            ptx::tc_fence_after();
            uint32_t r0[32];
            ptx::tmem_ld32(lane_addr + (uint32_t)(s * BKV + hf * 32), r0);
            ptx::tmem_ld_wait();
            const uint32_t km = (uint32_t)(kmask[s] >> (hf * 32));
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&s_free[s]);

            float v[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r0[i]) * sc;
            if (km != ~0u) {
#pragma unroll
                for (int i = 0; i < 32; ++i)
                    if (!((km >> i) & 1u)) v[i] = -INFINITY;
            }
            const bool band_edge = p.window >= 0 && ((q0 + BQ - 1) - k0 > p.window || (k0 + 31) - q0 > p.window);
            if (band_edge) {
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const int d = q - (k0 + i);
                    if (d > p.window || d < -p.window) v[i] = -INFINITY;
                }
            }
            float mx = v[0];
#pragma unroll
            for (int i = 1; i < 32; ++i) mx = fmaxf(mx, v[i]);
            // does ANY row of the CTA have to move its reference max at this tile?  (OR-reduction barrier, 256 threads)
            const int need_local = mx > m_used + RESCALE_THRESHOLD ? 1 : 0;
            int any;
            asm volatile(
                "{\n\t.reg .pred p, q;\n\t"
                "setp.ne.b32 p, %1, 0;\n\t"
                "barrier.red.or.pred q, 1, 256, p;\n\t"
                "selp.b32 %0, 1, 0, q;\n\t}"
                : "=r"(any)
                : "r"(need_local)
                : "memory");
            float alpha = 1.0f;
            if (any) {  // rare (always at j = 0): agree on the row max across the two halves, rescale O
                if (j > 0) {
                    ptx::mbar_wait(pv_done, (j - 1) & 1);  // P buffer idle, O quiescent
                    ptx::tc_fence_after();
                }
                xch[hf * 128 + r] = mx;
                asm volatile("bar.sync 1, 256;" ::: "memory");
                const float m_row = fmaxf(mx, xch[(hf ^ 1) * 128 + r]);
                const bool need = m_row > m_used + RESCALE_THRESHOLD;
                if (need) {
                    alpha = ptx::ex2(m_used - m_row);
                    m_used = m_row;
                }
                if (j > 0 && __any_sync(0xffffffffu, need)) {
#pragma unroll 1
                    for (int c = 0; c < 2; ++c) {  // this thread's half of the O columns
                        uint32_t o[32];
                        const uint32_t oa = lane_addr + 128u + (uint32_t)(hf * 64 + c * 32);
                        ptx::tmem_ld32(oa, o);
                        ptx::tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                        ptx::tmem_st32(oa, o);
                    }
                    ptx::tmem_st_wait();
                }
                asm volatile("bar.sync 1, 256;" ::: "memory");  // scratch read by everyone before P overwrites it
            }
            const float msafe = (m_used == -INFINITY) ? 0.f : m_used;
            uint32_t ph[16];
            float psum = 0.f;
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float p0 = ptx::ex2(v[2 * i] - msafe), p1 = ptx::ex2(v[2 * i + 1] - msafe);
                psum += p0 + p1;
                const __half2 hh = __floats2half2_rn(p0, p1);
                ph[i] = *reinterpret_cast<const uint32_t*>(&hh);
            }
            l = l * alpha + psum;
            if (j > 0 && !any) {
                ptx::mbar_wait(pv_done, (j - 1) & 1);
                ptx::tc_fence_after();
            }
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int cc = hf * 4 + c;
                *reinterpret_cast<uint4*>(sP + r * 128 + ((cc ^ (r & 7)) << 4)) = make_uint4(ph[4 * c], ph[4 * c + 1], ph[4 * c + 2], ph[4 * c + 3]);
            }
            ptx::fence_proxy_async_smem();
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(p_ready);
        }
        // ---- finalise: total row sum across the two halves, O / l -> fp16 (each thread writes 64 of the 128 columns)
        ptx::mbar_wait(pv_done, (n_tiles - 1) & 1);
        ptx::tc_fence_after();
        xch[hf * 128 + r] = l;
        asm volatile("bar.sync 1, 256;" ::: "memory");
        const float lt = l + xch[(hf ^ 1) * 128 + r];
        const float inv = lt > 0.f ? 1.0f / lt : 0.f;
        __half* op = p.O + ((long)b * p.q_len + q) * p.o_stride + (long)h * HD + hf * 64;
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
            uint32_t o[32];
            ptx::tmem_ld32(lane_addr + 128u + (uint32_t)(hf * 64 + c * 32), o);
            ptx::tmem_ld_wait();
            if (q < p.q_len) {
                uint32_t w[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const __half2 hh = __floats2half2_rn(__uint_as_float(o[2 * i]) * inv, __uint_as_float(o[2 * i + 1]) * inv);
                    w[i] = *reinterpret_cast<const uint32_t*>(&hh);
                }
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    *reinterpret_cast<uint4*>(op + c * 32 + i * 8) = make_uint4(w[4 * i], w[4 * i + 1], w[4 * i + 2], w[4 * i + 3]);
            }
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 9) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<TMEM_COLS>(tmem_base);
    }
}

// ---------------------------------------------------------------- P in tensor memory, two interleaved tile chains
// The 4-/8-warp kernels above run ONE dependency chain per CTA (S_j -> softmax_j -> P_j via smem -> PV_j -> pv_done ->
// P_{j+1} may be written): ~3.4 k cycles per key tile against 512 cycles of MMA (ncu: tensor 30 %, issue 31 %, MUFU 33 %).
// Here P_j is written back into the TMEM columns S_j came from (tcgen05.st, fp16 pairs) and the second MMA takes its A
// operand from tensor memory.  With S double-buffered that makes P double-buffered too, and the single-thread issue order
//     S_0 S_1 | PV_0 S_2 | PV_1 S_3 | ...         (tensor-core ops of one thread execute in order: S_{j+2} overwrites
//                                                  buffer j&1 only after PV_j has read P_j from it)
// leaves two independent chains (even / odd tiles) in flight: while the tensor core runs PV_j and S_{j+2}, the softmax warps
// are already on tile j+1.  No P buffer in shared memory (96 KB per CTA), no generic->async proxy fence, no wait on the
// previous PV before P is written; the only time the softmax warps wait for a PV is the rare lazy rescale of O.
constexpr int P3_OFF_K = Q_BYTES;
constexpr int P3_OFF_V = P3_OFF_K + 2 * (KV_TILE / 2) * 2;   // K stages: 2 x 16 KB
constexpr int P3_OFF_BAR = P3_OFF_V + 2 * KV_TILE;           // V stages: 2 x 16 KB  -> 98 304
constexpr int ATT_P3_SMEM = P3_OFF_BAR + 256;

__global__ void __launch_bounds__(ATT_TC_THREADS, 2)
attention_tcp_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                     const __grid_constant__ CUtensorMap tmV, const AttnParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sQ = smem;
    uint8_t* sK = smem + P3_OFF_K;
    uint8_t* sV = smem + P3_OFF_V;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + P3_OFF_BAR);
    uint64_t* q_full = bars;           // [1]
    uint64_t* k_full = bars + 1;       // [2]
    uint64_t* k_empty = bars + 3;      // [2]  S_j completed
    uint64_t* v_full = bars + 5;       // [2]
    uint64_t* v_empty = bars + 7;      // [2]  PV_j completed
    uint64_t* s_ready = bars + 9;      // [2]
    uint64_t* p_ready = bars + 11;     // [2]  4 softmax warps
    uint64_t* pv_done = bars + 13;     // [2]  by tile parity
    uint64_t* kmask = bars + 15;       // [4]  by tile index & 3
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 19);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q0 = blockIdx.x * BQ;
    const int h = blockIdx.y, b = blockIdx.z;
    const int hk = h / p.n_rep;

    int kb_lo = 0, kb_hi = (p.k_len + BKV - 1) / BKV - 1;
    if (p.window >= 0) {
        kb_lo = max(0, (q0 - p.window) / BKV);
        kb_hi = min(kb_hi, (min(q0 + BQ - 1, p.q_len - 1) + p.window) / BKV);
    }
    const int n_tiles = kb_hi - kb_lo + 1;

    if (threadIdx.x == 0) {
        if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();
        ptx::mbar_init(q_full, 1);
        for (int s = 0; s < 2; ++s) {
            ptx::mbar_init(&k_full[s], 1);
            ptx::mbar_init(&k_empty[s], 1);
            ptx::mbar_init(&v_full[s], 1);
            ptx::mbar_init(&v_empty[s], 1);
            ptx::mbar_init(&s_ready[s], 1);
            ptx::mbar_init(&p_ready[s], 4);
            ptx::mbar_init(&pv_done[s], 1);
        }
        ptx::fence_mbar_init();
    }
    if (warp == 5) ptx::tmem_alloc<TMEM_COLS>(tmem_slot);
    if (warp == 4 && lane == 0) {
        ptx::prefetch_tmap(&tmQ);
        ptx::prefetch_tmap(&tmK);
        ptx::prefetch_tmap(&tmV);
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    ptx::pdl_wait();
    ptx::pdl_launch();

    if (warp == 4) {
        // ===================== producer =====================
        if (ptx::elect_one()) {
            ptx::mbar_expect_tx(q_full, Q_BYTES);
            ptx::tma_load_4d(sQ, &tmQ, q_full, 0, h, q0, b);
            ptx::tma_load_4d(sQ + Q_BYTES / 2, &tmQ, q_full, 64, h, q0, b);
        }
        const int32_t* mb = p.key_mask ? p.key_mask + (long)b * p.k_len : nullptr;
        for (int j = 0; j < n_tiles; ++j) {
            const int s = j & 1;
            const uint32_t ph = (j >> 1) & 1;
            const int k0 = (kb_lo + j) * BKV;
            const int ka = k0 + lane, kb2 = k0 + 32 + lane;
            const bool ok_a = ka < p.k_len && (!mb || mb[ka] != 0);
            const bool ok_b = kb2 < p.k_len && (!mb || mb[kb2] != 0);
            const uint32_t lo = __ballot_sync(0xffffffffu, ok_a), hi = __ballot_sync(0xffffffffu, ok_b);
            ptx::mbar_wait(&k_empty[s], ph ^ 1);
            if (ptx::elect_one()) {
                kmask[j & 3] = (uint64_t)lo | ((uint64_t)hi << 32);
                ptx::mbar_expect_tx(&k_full[s], KV_TILE);
                ptx::tma_load_4d(sK + s * KV_TILE, &tmK, &k_full[s], 0, hk, k0, b);
                ptx::tma_load_4d(sK + s * KV_TILE + KV_TILE / 2, &tmK, &k_full[s], 64, hk, k0, b);
            }
            ptx::mbar_wait(&v_empty[s], ph ^ 1);
            if (ptx::elect_one()) {
                ptx::mbar_expect_tx(&v_full[s], KV_TILE);
                ptx::tma_load_4d(sV + s * KV_TILE, &tmV, &v_full[s], 0, hk, k0, b);
                ptx::tma_load_4d(sV + s * KV_TILE + KV_TILE / 2, &tmV, &v_full[s], 64, hk, k0, b);
            }
            __syncwarp();
        }
    } else if (warp == 5) {
        // ===================== MMA issuer =====================
        // whole warp on warp-uniform values, one elected lane issues: tcgen05.mma takes uniform-register operands, a
        // `lane == 0` block on per-thread values costs an ELECT + R2UR waterfall per instruction (see gemm_sm100.cu)
        constexpr uint32_t idesc_s = ptx::umma_idesc_f16(BQ, BKV);
        constexpr uint32_t idesc_o = ptx::umma_idesc_f16_bmn(BQ, HD);
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
        const uint64_t dq0 = ptx::umma_desc_sw128(ptx::smem_u32(sQ));
        const uint64_t dk0 = ptx::umma_desc_sw128(ptx::smem_u32(sK));
        const uint64_t dv0 = ptx::umma_desc_mn_sw128(ptx::smem_u32(sV), KV_TILE / 2, 1024);
        auto issue_s = [&](int j) {  // S[j&1] = Q . K_j^T, then release the K stage
            const int s = j & 1;
            const uint64_t dk = dk0 + (uint64_t)(s * (KV_TILE >> 4));
            if (ptx::elect_one()) {
#pragma unroll
                for (int kk = 0; kk < HD / 16; ++kk) {
                    const uint64_t da = dq0 + (uint64_t)(((kk >> 2) * (Q_BYTES / 2) + (kk & 3) * 32) >> 4);
                    const uint64_t db = dk + (uint64_t)(((kk >> 2) * (KV_TILE / 2) + (kk & 3) * 32) >> 4);
                    ptx::umma_f16(tmem_u + (uint32_t)(s * BKV), da, db, idesc_s, kk > 0 ? 1u : 0u);
                }
                ptx::umma_commit(&s_ready[s]);
                ptx::umma_commit(&k_empty[s]);
            }
            __syncwarp();
        };
        ptx::mbar_wait(q_full, 0);
        ptx::mbar_wait(&k_full[0], 0);
        ptx::tc_fence_after();
        issue_s(0);
        if (n_tiles > 1) {
            ptx::mbar_wait(&k_full[1], 0);
            ptx::tc_fence_after();
            issue_s(1);
        }
        for (int j = 0; j < n_tiles; ++j) {
            const int s = j & 1;
            const uint32_t ph = (j >> 1) & 1;
            ptx::mbar_wait(&p_ready[s], ph);
            ptx::mbar_wait(&v_full[s], ph);
            ptx::tc_fence_after();
            const uint64_t dv = dv0 + (uint64_t)(s * (KV_TILE >> 4));
            const uint32_t acc = j > 0 ? 1u : 0u;
            if (ptx::elect_one()) {
#pragma unroll
                for (int kk = 0; kk < BKV / 16; ++kk)  // A = P_j from TMEM: 16 keys = 8 packed columns per step
                    ptx::umma_f16_ts(tmem_u + 128u, tmem_u + (uint32_t)(s * BKV + kk * 8), dv + (uint64_t)((kk * 2048) >> 4), idesc_o, kk > 0 ? 1u : acc);
                ptx::umma_commit(&v_empty[s]);
                ptx::umma_commit(&pv_done[s]);
            }
            __syncwarp();
            if (j + 2 < n_tiles) {
                ptx::mbar_wait(&k_full[s], ((j + 2) >> 1) & 1);
                ptx::tc_fence_after();
                issue_s(j + 2);
            }
        }
    } else {
        // ===================== softmax (warps 0..3): thread = query row = TMEM lane =====================
        const int r = warp * 32 + lane;
        const int q = q0 + r;
        const uint32_t lane_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
        const float sc = 0.08838834764831845f * 1.4426950408889634f;
        float m_used = -INFINITY, l = 0.f;
        for (int j = 0; j < n_tiles; ++j) {
            const int s = j & 1;
            const int k0 = (kb_lo + j) * BKV;
            ptx::mbar_wait(&s_ready[s], (j >> 1) & 1);
            ptx::tc_fence_after();
            uint32_t r0[32], r1[32];
            ptx::tmem_ld32(lane_addr + (uint32_t)(s * BKV), r0);
            ptx::tmem_ld32(lane_addr + (uint32_t)(s * BKV + 32), r1);
            ptx::tmem_ld_wait();
            const uint64_t km = kmask[j & 3];
            float v[64];
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                v[i] = __uint_as_float(r0[i]) * sc;
                v[32 + i] = __uint_as_float(r1[i]) * sc;
            }
            const bool band_edge = p.window >= 0 && ((q0 + BQ - 1) - k0 > p.window || (k0 + BKV - 1) - q0 > p.window);
            if (km != ~0ull) {
                const uint32_t klo = (uint32_t)km, khi = (uint32_t)(km >> 32);
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    if (!((klo >> i) & 1u)) v[i] = -INFINITY;
                    if (!((khi >> i) & 1u)) v[32 + i] = -INFINITY;
                }
            }
            if (band_edge) {
#pragma unroll
                for (int i = 0; i < 64; ++i) {
                    const int d = q - (k0 + i);
                    if (d > p.window || d < -p.window) v[i] = -INFINITY;
                }
            }
            float mx0 = v[0], mx1 = v[1], mx2 = v[2], mx3 = v[3];  // four short chains instead of one 64-long
#pragma unroll
            for (int i = 4; i < 64; i += 4) {
                mx0 = fmaxf(mx0, v[i]); mx1 = fmaxf(mx1, v[i + 1]); mx2 = fmaxf(mx2, v[i + 2]); mx3 = fmaxf(mx3, v[i + 3]);
            }
            const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
            const float m_new = fmaxf(m_used, mx);
            const bool need = m_new > m_used + RESCALE_THRESHOLD;
            const float alpha = need ? ptx::ex2(m_used - m_new) : 1.0f;
            if (need) m_used = m_new;
            if (j > 0 && __any_sync(0xffffffffu, need)) {  // rare: O must be quiescent -> the previous PV has completed
                ptx::mbar_wait(&pv_done[(j - 1) & 1], ((j - 1) >> 1) & 1);
                ptx::tc_fence_after();
#pragma unroll 1
                for (int c = 0; c < 4; ++c) {
                    uint32_t o[32];
                    ptx::tmem_ld32(lane_addr + 128u + (uint32_t)(c * 32), o);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                    ptx::tmem_st32(lane_addr + 128u + (uint32_t)(c * 32), o);
                }
            }
            const float msafe = (m_used == -INFINITY) ? 0.f : m_used;
            uint32_t ph[32];
            float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const float p0 = ptx::ex2(v[2 * i] - msafe), p1 = ptx::ex2(v[2 * i + 1] - msafe);
                ps0 += p0;
                ps1 += p1;
                const __half2 hh = __floats2half2_rn(p0, p1);
                ph[i] = *reinterpret_cast<const uint32_t*>(&hh);
            }
            l = l * alpha + (ps0 + ps1);
            ptx::tmem_st32(lane_addr + (uint32_t)(s * BKV), ph);  // P_j over the first 32 columns of S_j
            ptx::tmem_st_wait();
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&p_ready[s]);
        }
        // ---- finalise
        ptx::mbar_wait(&pv_done[(n_tiles - 1) & 1], ((n_tiles - 1) >> 1) & 1);
        ptx::tc_fence_after();
        const float inv = l > 0.f ? 1.0f / l : 0.f;
        __half* op = p.O + ((long)b * p.q_len + q) * p.o_stride + (long)h * HD;
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
            uint32_t o[32];
            ptx::tmem_ld32(lane_addr + 128u + (uint32_t)(c * 32), o);
            ptx::tmem_ld_wait();
            if (q < p.q_len) {
                uint32_t w[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const __half2 hh = __floats2half2_rn(__uint_as_float(o[2 * i]) * inv, __uint_as_float(o[2 * i + 1]) * inv);
                    w[i] = *reinterpret_cast<const uint32_t*>(&hh);
                }
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    *reinterpret_cast<uint4*>(op + c * 32 + i * 8) = make_uint4(w[4 * i], w[4 * i + 1], w[4 * i + 2], w[4 * i + 3]);
            }
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 5) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<TMEM_COLS>(tmem_base);
    }
}

bool make_qkv_tmap(CUtensorMap* tm, const __half* base, long row_stride, int rows, int heads, int B, int box_rows) {
    const uint64_t dims[4] = {(uint64_t)HD, (uint64_t)heads, (uint64_t)rows, (uint64_t)B};
    const uint64_t strides[3] = {(uint64_t)HD * 2, (uint64_t)row_stride * 2, (uint64_t)rows * (uint64_t)row_stride * 2};
    const uint32_t box[4] = {64, 1, (uint32_t)box_rows, 1};
    return make_tmap_f16_sw128(tm, base, 4, dims, strides, box);
}

}  // namespace

bool attention_use_tc() {
    const char* e = std::getenv("ACE_B200_ATTN");  // read per call: the tests flip it between launches
    return !(e && std::string(e) == "mma");
}

cudaError_t launch_attention_tc(const __half* Q, long q_stride, const __half* K, long k_stride, const __half* V, long v_stride,
                                __half* O, long o_stride, const int32_t* key_mask, int B, int q_len, int k_len, int n_heads,
                                int n_kv_heads, int window, cudaStream_t st) {
    CUtensorMap tq, tk, tv;
    if (!make_qkv_tmap(&tq, Q, q_stride, q_len, n_heads, B, BQ) || !make_qkv_tmap(&tk, K, k_stride, k_len, n_kv_heads, B, BKV) ||
        !make_qkv_tmap(&tv, V, v_stride, k_len, n_kv_heads, B, BKV))
        return cudaErrorInvalidValue;
    static DeviceOnce attr;
    const cudaError_t ea = attr.run([](int&) {
        cudaError_t e = cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_TC_SMEM);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(attention_tc8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_TC_SMEM);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(attention_tcp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ATT_P3_SMEM);
        return e;
    });
    if (ea != cudaSuccess) return ea;
    AttnParams p{O, o_stride, key_mask, q_len, k_len, n_heads / n_kv_heads, window};
    dim3 grid((q_len + BQ - 1) / BQ, n_heads, B);
    const char* env = std::getenv("ACE_B200_ATTN");
    if (env && std::string(env) == "tc4") return launch_pdl(attention_tc_kernel, grid, dim3(ATT_TC_THREADS), (size_t)ATT_TC_SMEM, st, tq, tk, tv, p);
    if (env && std::string(env) == "tc8") return launch_pdl(attention_tc8_kernel, grid, dim3(ATT8_THREADS), (size_t)ATT_TC_SMEM, st, tq, tk, tv, p);
    return launch_pdl(attention_tcp_kernel, grid, dim3(ATT_TC_THREADS), (size_t)ATT_P3_SMEM, st, tq, tk, tv, p);
}

}  // namespace ace
