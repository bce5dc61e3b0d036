// One Oobleck residual unit at 128 channels in ONE launch (sm_100a): the two last decoder levels, 63 % of a VAE decode.
//
//     h   = snake2( conv_k7_dilated(xs) + b1 )         tap-GEMM, 14 k-blocks, A from an activation slab (as gemm_tc_kernel<SLAB>)
//     x'  = x + conv_k1(h) + b2                         second MMA: h goes registers -> shared memory -> tensor core, never to HBM
//     out_x  = fp16(x')        (the 16-bit residual stream, optional)
//     out_xs = fp16(snake_next(x'))                     (input of the next unit / block)
// reference: acestep_vae_model.cpp:682-742 (snake, dilated conv, k1 conv, residual add of one unit)
//
// Unfused, the k1 convolution is a pure HBM stream (h and x in, x' and xs out: 8 bytes per element at 128 channels against 256 FLOPs)
// running at 0.49 of the copy bandwidth (profiles/r2_vae240_launches_p.md) — 36 % of the decode.  Fused, h (a quarter of that
// traffic plus the k7 kernel's store) disappears and the remaining three streams ride under the k7 MMAs.
// The arithmetic is the unfused kernels' to the bit: same MMA order, h rounded to fp16 before the second MMA, (acc + bias) + x in fp32.
//
// CTA = 10 warps, persistent; by default a CLUSTER OF TWO CTAs (cta_group::2, M = 256) works on 256-row items, CTA r owning rows 128 r..
// (template PAIR; the single-CTA variant on 128-row tiles is kept behind ACE_B200_VAE_FUSE=1):
//   warps 0-7  epilogue: warp w owns TMEM lanes 32 (w & 3).. and the 64-column half (w >> 2) of both accumulators
//   warp  8    TMA producer: activation slab per item (double-buffered), weight ring (slot = the two 64-channel k-blocks of a tap, a pair
//              CTA stages half of each; the k1 weights travel through the same ring, one slot per item)
//   warp  9    TMEM allocator + tcgen05.mma issuer (leader CTA; CTA 1's warp 9 relays its "landed" barriers to the leader).  Issue order
//              MMA1(0) | MMA1(1) MMA2(0) | MMA1(2) MMA2(1) | ... : the second MMA of an item is issued one item late, so the tensor pipe
//              never waits for the epilogue pass that produces h.
// h: pair — pass 1 writes it as fp16 pairs into TENSOR MEMORY (columns 384.., double-buffered by item parity) and the second MMA takes its
// A operand from there; single CTA — a K-major SWIZZLE_128B image in shared memory, whose 4 KB per warp doubles as the staging slice of
// the {64, 32} TMA store boxes that carry x' and xs out.  Epilogue order per iteration i (pair): pass 1 of item i, then pass 2 of item
// i - 1 (its second MMA was issued an iteration ago); the "h ready" barriers are therefore per item parity (see below).
#include "vae.h"
#include "gemm.h"
#include "kernels.h"
#include "ptx.cuh"
#include "tmap.h"

#include <cuda.h>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

namespace ace {

namespace {

constexpr int RU_C = 128;                      // channels in = out
constexpr int RU_BM = 128;                     // rows per tile
constexpr int RU_TAPS = 7;
constexpr int RU_THREADS = 320;
constexpr int RU_SLAB_MAX_ROWS = 184;          // 128 + 6 x dilation 9 = 182 -> whole 8-row swizzle atoms
constexpr int RU_PANEL = RU_SLAB_MAX_ROWS * 128;   // one 64-channel panel of a slab (23 KB)
constexpr int RU_SLAB = 2 * RU_PANEL;
constexpr int RU_WTILE = 128 * 64 * 2;         // one packed F16 weight tile: 128 output channels x 64 input channels (16 KB)
constexpr int RU_HTILE = 2 * RU_BM * 128;      // h as MMA operand: two 64-channel panels of 128 rows (32 KB)
constexpr int RU_COLS = 6 * RU_C * 4;          // bias1, snake2 a, snake2 1/b, bias2, next a, next 1/b
// PAIR = a cluster of two CTAs on cta_group::2 MMAs (M = 256 across the pair): CTA r owns rows l0 + 128 r of a 256-row super tile
// (its own slab, accumulators, h and stores) and stages only HALF of every weight tile (output channels 64 r ..), the tensor core reads
// the other half from the peer's shared memory.  The single-CTA kernel re-reads 229 KB of weights per 128-row tile and sits at the L2
// throughput cap (ncu: 29.6 GB L2 -> SM per launch at 3.9 ms = 7.6 TB/s; ~330 KB per tile against 6300 B/cycle chip-wide);
// the pair halves the weight bytes per row, and the shared memory it frees holds a separate staging buffer for the stores.
template <bool PAIR>
struct RuLayout {
    static constexpr int WHALF = PAIR ? RU_WTILE / 2 : RU_WTILE;   // bytes of one k-block's weights staged by this CTA
    static constexpr int SLOT = 2 * WHALF;                         // ring slot = both k-blocks of a tap
    static constexpr int NSLOT = PAIR ? 4 : 3;
    static constexpr int OFF_W = 2 * RU_SLAB;
    static constexpr int OFF_H = OFF_W + NSLOT * SLOT;
    static constexpr int OFF_STG = OFF_H + RU_HTILE;               // PAIR: landing buffer of the residual boxes (32 KB)
    static constexpr int OFF_COL = OFF_STG + (PAIR ? RU_HTILE : 0);
    static constexpr int OFF_BAR = OFF_COL + RU_COLS;
    static constexpr int SMEM = OFF_BAR + 512;
    static_assert(SMEM <= 227 * 1024, "residual-unit kernel exceeds shared memory");
    static_assert(RU_PANEL % 1024 == 0 && OFF_W % 1024 == 0 && OFF_H % 1024 == 0 && OFF_STG % 1024 == 0, "swizzled tiles need 1 KB alignment");
};

struct RuParams {
    const uint8_t* W1;       // 14 packed tiles (tap-major, two k-blocks per tap)
    const uint8_t* W2;       // 2 packed tiles
    const float* cols[6];    // bias1, snake2 alpha, snake2 1/beta, bias2, next alpha, next 1/beta   [128] each
    const __half* x_in;      // residual stream [B, L, 128]
    int B, L, m_tiles_per_b, dilation, slab_rows, store_x;
    long long* trace;        // debug (ACE_B200_RU_TRACE=1): clock64 stamps of one steady-state tile per CTA, else null
    int stall_ns;            // test hook (ACE_B200_RU_STALL_NS): the producer sleeps this long before every W2 load, so the second MMA
                             // of an item is issued late and the epilogue warps run ahead of the issuer (see h_ready below)
};

constexpr int RU_TR_N = 32, RU_TR_TILE = 4;   // stamps per CTA; the CTA's 5th work item is the one traced
__device__ __forceinline__ void ru_stamp(const RuParams& p, int it, int slot) {
    if (p.trace && it == RU_TR_TILE) p.trace[(size_t)blockIdx.x * RU_TR_N + slot] = clock64();
}

__device__ __forceinline__ float snake_f(float v, float a, float inv_b) {  // gemm_sm100.cu's, to the instruction
    const float s = __sinf(a * v);
    return fmaf(s * s, inv_b, v);
}
__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
    const __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&h);
}

template <bool PAIR>
__global__ void __launch_bounds__(RU_THREADS, 1)
vae_resunit128_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmX,
                      const __grid_constant__ CUtensorMap tmXS, const __grid_constant__ CUtensorMap tmR, const RuParams p) {
    using LY = RuLayout<PAIR>;
    constexpr int NSLOT = LY::NSLOT;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sSlab = smem;                     // [2][2 panels][184 rows x 128 B]
    uint8_t* sW = smem + LY::OFF_W;            // [NSLOT][2 k-blocks x WHALF]
    uint8_t* sH = smem + LY::OFF_H;            // [2 panels][128 rows x 128 B]: single CTA: h, then x' / xs on their way out; PAIR (h lives in tensor memory): x' / xs only
    uint8_t* sStg = smem + LY::OFF_STG;        // PAIR: residual rows as they land, same geometry as sH
    float* sCol = reinterpret_cast<float*>(smem + LY::OFF_COL);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + LY::OFF_BAR);
    uint64_t* full = bars;              // [4] weight slot landed (local)
    uint64_t* empty = bars + 4;         // [4] MMAs reading the slot have completed (PAIR: multicast commit)
    uint64_t* peer_full = bars + 8;     // [4] leader only: CTA 1's slot landed
    uint64_t* slab_full = bars + 12;    // [2]
    uint64_t* slab_empty = bars + 14;   // [2]
    uint64_t* peer_slab = bars + 16;    // [2] leader only
    uint64_t* acc1_full = bars + 18;    // [2]
    uint64_t* acc1_empty = bars + 20;   // [2] leader only in a pair: the epilogue warps of both CTAs
    // h_ready / peer_h are indexed by item parity and waited on with parity (item >> 1) & 1.  In a pair, pass 1 of item j + 1 does not
    // depend on the second MMA of item j, so the epilogue warps can finish it while the issuer still waits for item j's W2 slot; a single
    // barrier waited on with parity j & 1 would then be TWO phases ahead of the issuer, the wait would never succeed and the kernel hung
    // (seen once in ~20 decodes of a 240 s clip).  Per-parity barriers cannot lead by more than one phase: pass 1 of item j + 2 needs the
    // first MMA of item j + 2, which the issuer starts after the second MMA of item j.
    uint64_t* h_ready = bars + 36;      // [2] 8 local epilogue warps: h of the item is in shared / tensor memory
    uint64_t* peer_h = bars + 38;       // [2] leader only: CTA 1's h
    uint64_t* acc2_full = bars + 24;
    uint64_t* acc2_empty = bars + 25;   //     leader only in a pair
    uint64_t* r_full = bars + 26;       // [8] PAIR: the residual box of epilogue warp w has landed in its staging slice
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 34);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = PAIR ? ptx::cluster_ctarank() : 0u;
    // work items: 128-row tiles (single CTA) or 256-row super tiles (pair; this CTA's rows start at 128 * rank)
    const int first = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
    const int stride = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    const int items_per_b = PAIR ? (p.m_tiles_per_b + 1) / 2 : p.m_tiles_per_b;
    const int total_tiles = p.B * items_per_b;
    constexpr int ITEM_ROWS = PAIR ? 2 * RU_BM : RU_BM;
    constexpr int EPI_ARRIVALS = PAIR ? 16 : 8;

    if (threadIdx.x == 0) {
        if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();
        for (int s = 0; s < NSLOT; ++s) {
            ptx::mbar_init(&full[s], 1);
            ptx::mbar_init(&empty[s], 1);
            ptx::mbar_init(&peer_full[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            ptx::mbar_init(&slab_full[a], 1);
            ptx::mbar_init(&slab_empty[a], 1);
            ptx::mbar_init(&peer_slab[a], 1);
            ptx::mbar_init(&acc1_full[a], 1);
            ptx::mbar_init(&acc1_empty[a], EPI_ARRIVALS);
        }
        for (int a = 0; a < 2; ++a) {
            ptx::mbar_init(&h_ready[a], 8);
            ptx::mbar_init(&peer_h[a], 1);
        }
        ptx::mbar_init(acc2_full, 1);
        ptx::mbar_init(acc2_empty, EPI_ARRIVALS);
        for (int w = 0; w < 8; ++w) ptx::mbar_init(&r_full[w], 1);
        ptx::fence_mbar_init();
    }
    if (warp == 9) {
        if (PAIR) ptx::tmem_alloc_2cta<512>(tmem_slot);
        else ptx::tmem_alloc<512>(tmem_slot);
    }
    if (warp == 8 && lane == 0) {
        ptx::prefetch_tmap(&tmA);
        ptx::prefetch_tmap(&tmX);
        ptx::prefetch_tmap(&tmXS);
        if (PAIR) ptx::prefetch_tmap(&tmR);
    }
    if (threadIdx.x < RU_C) {  // per-channel constants (weights: not produced by the previous kernel, safe before pdl_wait)
#pragma unroll
        for (int k = 0; k < 6; ++k) sCol[k * RU_C + threadIdx.x] = __ldg(p.cols[k] + threadIdx.x);
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (PAIR) ptx::cluster_sync();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    ptx::pdl_wait();
    ptx::pdl_launch();

    auto item_coords = [&](int t, int& b, int& l0) {
        b = t / items_per_b;
        l0 = (t % items_per_b) * ITEM_ROWS + (int)rank * RU_BM;
    };

    if (warp == 8) {
        // ===================== TMA producer (both CTAs of a pair) =====================
        uint32_t s = 0, ph = 0;
        auto ring_load = [&](const uint8_t* src) {   // src: the tap's two 16 KB tiles; a pair CTA takes rows 64 r.. of each (8 KB)
            ptx::mbar_wait(&empty[s], ph ^ 1);
            if (ptx::elect_one()) {
                ptx::mbar_expect_tx(&full[s], LY::SLOT);
                ptx::bulk_load(sW + s * LY::SLOT, src + rank * LY::WHALF, LY::WHALF, &full[s]);
                ptx::bulk_load(sW + s * LY::SLOT + LY::WHALF, src + RU_WTILE + rank * LY::WHALF, LY::WHALF, &full[s]);
            }
            __syncwarp();
            if (++s == (uint32_t)NSLOT) {
                s = 0;
                ph ^= 1;
            }
        };
        const uint32_t panel_bytes = (uint32_t)p.slab_rows * 128u;
        auto slab_load = [&](int i, int t) {   // activation slab of the CTA's i-th item
            int b, l0;
            item_coords(t, b, l0);
            const int buf = i & 1;
            ptx::mbar_wait(&slab_empty[buf], ((i >> 1) & 1) ^ 1);
            if (ptx::elect_one()) {
                ptx::mbar_expect_tx(&slab_full[buf], 2 * panel_bytes);
                ptx::tma_load_3d(sSlab + buf * RU_SLAB, &tmA, &slab_full[buf], 0, l0 - 3 * p.dilation, b);
                ptx::tma_load_3d(sSlab + buf * RU_SLAB + RU_PANEL, &tmA, &slab_full[buf], 64, l0 - 3 * p.dilation, b);
            }
            __syncwarp();
        };
        // The slab of item i + 1 is requested in the MIDDLE of item i's weight loads: the first NSLOT of those run ahead of the tensor
        // core, the next one waits for the first tap of item i to be consumed — by then MMA1(i - 1) has released the slab buffer.
        // Requested after ALL of item i's weights (program order of one thread) the slab arrived 1-2 k cycles late (trace).
        int it = 0;
        if (first < total_tiles) slab_load(0, first);
        for (int t = first; t < total_tiles; t += stride, ++it) {
            for (int tap = 0; tap < RU_TAPS; ++tap) {
                ring_load(p.W1 + (size_t)tap * (2 * RU_WTILE));
                if (tap == NSLOT && t + stride < total_tiles) slab_load(it + 1, t + stride);
            }
            if (it > 0) {   // consumed by MMA2 of the previous tile, issued after this tile's MMA1
                if (p.stall_ns) __nanosleep((unsigned)p.stall_ns);
                ring_load(p.W2);
            }
        }
        if (it > 0) {
            if (p.stall_ns) __nanosleep((unsigned)p.stall_ns);
            ring_load(p.W2);
        }
    } else if (warp == 9 && rank == 0) {
        // ===================== MMA issuer (leader; whole warp on uniform values, one elected lane issues) =====================
        constexpr uint32_t idesc = ptx::umma_idesc_f16(PAIR ? 2 * RU_BM : RU_BM, RU_C);
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
        const uint64_t desc_slab0 = ptx::umma_desc_sw128(ptx::smem_u32(sSlab));
        const uint64_t desc_w0 = ptx::umma_desc_sw128(ptx::smem_u32(sW));
        const uint64_t desc_h0 = ptx::umma_desc_sw128(ptx::smem_u32(sH));
        auto mma = [&](uint32_t d, uint64_t da, uint64_t db, uint32_t acc) {
            if (PAIR) ptx::umma_f16_2cta(d, da, db, idesc, acc);
            else ptx::umma_f16(d, da, db, idesc, acc);
        };
        auto commit = [&](uint64_t* bar) {   // PAIR: arrives on the barrier at this offset in BOTH CTAs
            if (PAIR) ptx::umma_commit_2cta(bar, 3);
            else ptx::umma_commit(bar);
        };
        uint32_t s = 0, ph = 0;
        auto ring_next = [&]() {
            if (++s == (uint32_t)NSLOT) {
                s = 0;
                ph ^= 1;
            }
        };
        auto mma2 = [&](int j) {  // acc2 = h_j . W2^T
            ptx::mbar_wait(&full[s], ph);
            if (PAIR) ptx::mbar_wait(&peer_full[s], ph);
            if (lane == 0) ru_stamp(p, j + 1, 4);
            ptx::mbar_wait(&h_ready[j & 1], (uint32_t)((j >> 1) & 1));
            if (PAIR) ptx::mbar_wait(&peer_h[j & 1], (uint32_t)((j >> 1) & 1));
            if (lane == 0) ru_stamp(p, j + 1, 5);
            ptx::mbar_wait(acc2_empty, (uint32_t)((j & 1) ^ 1));
            ptx::tc_fence_after();
            const uint64_t db = desc_w0 + (uint64_t)(s * (LY::SLOT >> 4));
            if (ptx::elect_one()) {
#pragma unroll
                for (int kc = 0; kc < 2; ++kc)
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint64_t dbk = db + (uint64_t)(kc * (LY::WHALF >> 4) + 2 * k);
                        if (PAIR)   // A = h_j from tensor memory: 16 channels = 8 packed columns per step
                            ptx::umma_f16_ts_2cta(tmem_u + 256u, tmem_u + 384u + (uint32_t)((j & 1) * 64 + (kc * 4 + k) * 8), dbk, idesc, (kc | k) ? 1u : 0u);
                        else
                            mma(tmem_u + 256u, desc_h0 + (uint64_t)(kc * ((RU_BM * 128) >> 4) + 2 * k), dbk, (kc | k) ? 1u : 0u);
                    }
                commit(&empty[s]);
                commit(acc2_full);
            }
            __syncwarp();
            ring_next();
        };
        int it = 0;
        for (int t = first; t < total_tiles; t += stride, ++it) {
            const int a = it & 1;
            if (lane == 0) ru_stamp(p, it, 0);
            ptx::mbar_wait(&acc1_empty[a], (uint32_t)(((it >> 1) & 1) ^ 1));
            if (lane == 0) ru_stamp(p, it, 1);
            ptx::mbar_wait(&slab_full[a], (uint32_t)((it >> 1) & 1));
            if (PAIR) ptx::mbar_wait(&peer_slab[a], (uint32_t)((it >> 1) & 1));
            ptx::tc_fence_after();
            if (lane == 0) ru_stamp(p, it, 2);
            const uint32_t d_tmem = tmem_u + (uint32_t)(a * RU_C);
            const uint64_t da0 = desc_slab0 + (uint64_t)(a * (RU_SLAB >> 4));
            for (int tap = 0; tap < RU_TAPS; ++tap) {
                ptx::mbar_wait(&full[s], ph);
                if (PAIR) ptx::mbar_wait(&peer_full[s], ph);
                ptx::tc_fence_after();
                const uint64_t da = da0 + (uint64_t)(tap * p.dilation * 8);   // `dilation` rows of 128 bytes per tap
                const uint64_t db = desc_w0 + (uint64_t)(s * (LY::SLOT >> 4));
                if (ptx::elect_one()) {
#pragma unroll
                    for (int kc = 0; kc < 2; ++kc)
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            mma(d_tmem, da + (uint64_t)(kc * (RU_PANEL >> 4) + 2 * k), db + (uint64_t)(kc * (LY::WHALF >> 4) + 2 * k), (tap | kc | k) ? 1u : 0u);
                    commit(&empty[s]);
                }
                __syncwarp();
                ring_next();
            }
            if (ptx::elect_one()) {
                commit(&acc1_full[a]);
                commit(&slab_empty[a]);
            }
            __syncwarp();
            if (lane == 0) ru_stamp(p, it, 3);
            if (it > 0) mma2(it - 1);
            if (lane == 0) ru_stamp(p, it, 6);
        }
        if (it > 0) mma2(it - 1);
    } else if (warp == 9) {
        // ===================== relay (CTA 1 of a pair): local "landed / written" -> the leader's peer barriers, in the leader's order ====
        if (lane == 0) {
            uint32_t s = 0, ph = 0;
            auto relay_slot = [&]() {
                ptx::mbar_wait(&full[s], ph);
                ptx::mbar_arrive_cluster(ptx::mapa(&peer_full[s], 0));
                if (++s == (uint32_t)NSLOT) {
                    s = 0;
                    ph ^= 1;
                }
            };
            auto relay_h = [&](int j) {
                relay_slot();   // the W2 slot
                ptx::mbar_wait(&h_ready[j & 1], (uint32_t)((j >> 1) & 1));
                ptx::mbar_arrive_cluster(ptx::mapa(&peer_h[j & 1], 0));
            };
            int it = 0;
            for (int t = first; t < total_tiles; t += stride, ++it) {
                const int a = it & 1;
                ptx::mbar_wait(&slab_full[a], (uint32_t)((it >> 1) & 1));
                ptx::mbar_arrive_cluster(ptx::mapa(&peer_slab[a], 0));
                for (int tap = 0; tap < RU_TAPS; ++tap) relay_slot();
                if (it > 0) relay_h(it - 1);
            }
            if (it > 0) relay_h(it - 1);
        }
    } else {
        // ===================== epilogue: warps 0..7 =====================
        const int q = warp & 3, hf = warp >> 2;
        const int r = q * 32 + lane;
        const int slice_off = hf * (RU_BM * 128) + q * (32 * 128);   // this warp's 32 rows x 64 channels: 4 KB, 1 KB aligned
        uint8_t* slice = sH + slice_off;                             // single CTA: h, then x' / xs on their way out; pair: x' / xs only
        uint8_t* my_row = slice + lane * 128;
        uint8_t* r_slice = sStg + slice_off;                         // pair: landing buffer of the residual box
        const uint8_t* r_row = r_slice + lane * 128;
        const int sw = lane & 7;
        auto arrive_leader = [&](uint64_t* bar) {   // barriers the MMA issuer waits on live in the leader CTA
            if (PAIR) ptx::mbar_arrive_cluster(ptx::mapa(bar, 0));
            else ptx::mbar_arrive(bar);
        };
        const uint32_t lane_taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(hf * 64);
        const float* cB1 = sCol + 0 * RU_C + hf * 64;
        const float* cA2 = sCol + 1 * RU_C + hf * 64;
        const float* cI2 = sCol + 2 * RU_C + hf * 64;
        const float* cB2 = sCol + 3 * RU_C + hf * 64;
        const float* cAn = sCol + 4 * RU_C + hf * 64;
        const float* cIn = sCol + 5 * RU_C + hf * 64;
        uint4 R[8];   // residual x of the tile in flight: this thread's row, 64 channels

        const bool tr = warp == 0 && lane == 0;
        // residual box of an item's rows into this warp's landing slice (pair).  Per-thread global loads of a row each (8 x LDG.128 at a
        // 256-byte lane stride) cost 32 LSU passes per instruction, ~2 k cycles per tile for the eight warps (trace); the box costs one
        // TMA instruction and 8 conflict-free LDS.128.  It is requested a whole iteration before it is read: under the producer's
        // slab / weight traffic a 4 KB box took ~2.5 k cycles to land.
        auto issue_residual = [&](int b, int l0) {
            if (lane == 0) {
                ptx::mbar_expect_tx(&r_full[warp], 32 * 128);
                ptx::tma_load_3d(r_slice, &tmR, &r_full[warp], hf * 64, l0 + q * 32, b);
            }
        };
        // ---- pass 2 of item j: acc2 -> x' = (acc + b2) + x -> fp16 out; xs = snake_next(x') -> fp16 out.  x' is staged and handed to the
        // TMA first, the Snake of the 64 values (the expensive part) runs while the TMA reads the slice, then xs re-uses the slice.
        // `fetch_next`: request the residual rows of item j + 1 (coordinates nb, nl0) as soon as this item's have been read.
        auto pass2 = [&](int j, int b, int l0, bool fetch_next, int nb, int nl0) {
            if (tr) ru_stamp(p, j + 1, 8);
            ptx::mbar_wait(acc2_full, (uint32_t)(j & 1));
            ptx::tc_fence_after();
            if (tr) ru_stamp(p, j + 1, 9);
            uint32_t v[2][32];
            ptx::tmem_ld32(lane_taddr + 256u, v[0]);
            ptx::tmem_ld32(lane_taddr + 256u + 32u, v[1]);
            ptx::tmem_ld_wait();
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) arrive_leader(acc2_empty);
            if (tr) ru_stamp(p, j + 1, 10);
            const bool rows_in = l0 + q * 32 < p.L;   // warp-uniform: a box entirely below the tensor is not issued
            if (PAIR) {
                ptx::mbar_wait(&r_full[warp], (uint32_t)(j & 1));
#pragma unroll
                for (int g = 0; g < 8; ++g) R[g] = *reinterpret_cast<const uint4*>(r_row + ((g ^ sw) << 4));
                __syncwarp();   // every lane has read its row
                if (fetch_next) issue_residual(nb, nl0);
                if (lane == 0) ptx::bulk_wait_read0();   // the previous item's xs store has read the slice (issued an iteration ago)
                __syncwarp();
                if (tr) ru_stamp(p, j + 1, 13);
            }
#pragma unroll
            for (int g = 0; g < 8; ++g) {   // 16-byte group g: channels 8 g .. 8 g + 7 of this warp's half
                uint32_t* vv = &v[g >> 2][(g & 3) * 8];
                const float4 b0 = *reinterpret_cast<const float4*>(cB2 + 8 * g), b1 = *reinterpret_cast<const float4*>(cB2 + 8 * g + 4);
                const float2 r0 = __half22float2(*reinterpret_cast<const __half2*>(&R[g].x)), r1 = __half22float2(*reinterpret_cast<const __half2*>(&R[g].y));
                const float2 r2 = __half22float2(*reinterpret_cast<const __half2*>(&R[g].z)), r3 = __half22float2(*reinterpret_cast<const __half2*>(&R[g].w));
                const float x0 = (__uint_as_float(vv[0]) + b0.x) + r0.x, x1 = (__uint_as_float(vv[1]) + b0.y) + r0.y;
                const float x2 = (__uint_as_float(vv[2]) + b0.z) + r1.x, x3 = (__uint_as_float(vv[3]) + b0.w) + r1.y;
                const float x4 = (__uint_as_float(vv[4]) + b1.x) + r2.x, x5 = (__uint_as_float(vv[5]) + b1.y) + r2.y;
                const float x6 = (__uint_as_float(vv[6]) + b1.z) + r3.x, x7 = (__uint_as_float(vv[7]) + b1.w) + r3.y;
                if (p.store_x)
                    *reinterpret_cast<uint4*>(my_row + ((g ^ sw) << 4)) = make_uint4(pack_h2(x0, x1), pack_h2(x2, x3), pack_h2(x4, x5), pack_h2(x6, x7));
                vv[0] = __float_as_uint(x0); vv[1] = __float_as_uint(x1); vv[2] = __float_as_uint(x2); vv[3] = __float_as_uint(x3);
                vv[4] = __float_as_uint(x4); vv[5] = __float_as_uint(x5); vv[6] = __float_as_uint(x6); vv[7] = __float_as_uint(x7);
            }
            if (p.store_x) {
                ptx::fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0 && rows_in) {
                    ptx::tma_store_3d(&tmX, slice, hf * 64, l0 + q * 32, b);
                    ptx::bulk_commit();
                }
            }
            if (tr) ru_stamp(p, j + 1, 11);
            uint4 XS[8];
#pragma unroll
            for (int g = 0; g < 8; ++g) {
                const uint32_t* vv = &v[g >> 2][(g & 3) * 8];
                const float4 a0 = *reinterpret_cast<const float4*>(cAn + 8 * g), a1 = *reinterpret_cast<const float4*>(cAn + 8 * g + 4);
                const float4 i0 = *reinterpret_cast<const float4*>(cIn + 8 * g), i1 = *reinterpret_cast<const float4*>(cIn + 8 * g + 4);
                XS[g] = make_uint4(pack_h2(snake_f(__uint_as_float(vv[0]), a0.x, i0.x), snake_f(__uint_as_float(vv[1]), a0.y, i0.y)),
                                   pack_h2(snake_f(__uint_as_float(vv[2]), a0.z, i0.z), snake_f(__uint_as_float(vv[3]), a0.w, i0.w)),
                                   pack_h2(snake_f(__uint_as_float(vv[4]), a1.x, i1.x), snake_f(__uint_as_float(vv[5]), a1.y, i1.y)),
                                   pack_h2(snake_f(__uint_as_float(vv[6]), a1.z, i1.z), snake_f(__uint_as_float(vv[7]), a1.w, i1.w)));
            }
            if (tr) ru_stamp(p, j + 1, 12);
            if (p.store_x) {
                if (lane == 0) ptx::bulk_wait_read0();   // the x' store has read the slice: the same 4 KB carries xs next
                __syncwarp();
            }
#pragma unroll
            for (int g = 0; g < 8; ++g) *reinterpret_cast<uint4*>(my_row + ((g ^ sw) << 4)) = XS[g];
            ptx::fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0 && rows_in) {
                ptx::tma_store_3d(&tmXS, slice, hf * 64, l0 + q * 32, b);
                ptx::bulk_commit();
            }
            if (tr) ru_stamp(p, j + 1, 14);
        };

        // ---- pass 1: acc1 -> h = snake2(acc + b1), the A operand of the second MMA.  Single CTA: K-major image in shared memory.
        // Pair: fp16 pairs in tensor memory (64 columns, double-buffered by item parity), so the pass neither waits for the second MMA
        // of the previous item to release a buffer nor needs a proxy fence.
        auto pass1 = [&](int i) {
            const int a = i & 1;
            if (tr) ru_stamp(p, i, 15);
            ptx::mbar_wait(&acc1_full[a], (uint32_t)((i >> 1) & 1));
            ptx::tc_fence_after();
            if (tr) ru_stamp(p, i, 16);
            uint32_t HP[32];   // h of this thread's row, packed: 64 channels
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                uint32_t v[32];
                ptx::tmem_ld32(lane_taddr + (uint32_t)(a * RU_C + c * 32), v);
                ptx::tmem_ld_wait();
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    const int ch = c * 32 + g * 8;
                    const float4 b0 = *reinterpret_cast<const float4*>(cB1 + ch), b1 = *reinterpret_cast<const float4*>(cB1 + ch + 4);
                    const float4 a0 = *reinterpret_cast<const float4*>(cA2 + ch), a1 = *reinterpret_cast<const float4*>(cA2 + ch + 4);
                    const float4 i0 = *reinterpret_cast<const float4*>(cI2 + ch), i1 = *reinterpret_cast<const float4*>(cI2 + ch + 4);
                    const uint32_t* vv = &v[g * 8];
                    uint32_t* o = &HP[(c * 4 + g) * 4];
                    o[0] = pack_h2(snake_f(__uint_as_float(vv[0]) + b0.x, a0.x, i0.x), snake_f(__uint_as_float(vv[1]) + b0.y, a0.y, i0.y));
                    o[1] = pack_h2(snake_f(__uint_as_float(vv[2]) + b0.z, a0.z, i0.z), snake_f(__uint_as_float(vv[3]) + b0.w, a0.w, i0.w));
                    o[2] = pack_h2(snake_f(__uint_as_float(vv[4]) + b1.x, a1.x, i1.x), snake_f(__uint_as_float(vv[5]) + b1.y, a1.y, i1.y));
                    o[3] = pack_h2(snake_f(__uint_as_float(vv[6]) + b1.z, a1.z, i1.z), snake_f(__uint_as_float(vv[7]) + b1.w, a1.w, i1.w));
                }
            }
            if (tr) ru_stamp(p, i, 17);
            if (PAIR) {
                ptx::tmem_st32(tmem_base + ((uint32_t)(q * 32) << 16) + 384u + (uint32_t)(a * 64 + hf * 32), HP);
                ptx::tmem_st_wait();
            } else {
                // the previous tile's xs store has read this warp's 4 KB (the h slice doubles as store staging)
                if (lane == 0) ptx::bulk_wait_read0();
                __syncwarp();
#pragma unroll
                for (int g = 0; g < 8; ++g)
                    *reinterpret_cast<uint4*>(my_row + ((g ^ sw) << 4)) = make_uint4(HP[4 * g], HP[4 * g + 1], HP[4 * g + 2], HP[4 * g + 3]);
                ptx::fence_proxy_async_smem();
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                arrive_leader(&acc1_empty[a]);
                ptx::mbar_arrive(&h_ready[a]);   // local: the leader's issuer waits on it directly, CTA 1's relay forwards it
            }
            if (tr) ru_stamp(p, i, 18);
        };

        int it = 0, pb = 0, pl0 = 0;
        if (PAIR && first < total_tiles) {
            int b, l0;
            item_coords(first, b, l0);
            issue_residual(b, l0);
        }
        for (int t = first; t < total_tiles; t += stride, ++it) {
            int b, l0;
            item_coords(t, b, l0);
            if (PAIR) {
                // pass 1 of this item first, then pass 2 of the previous one: the second MMA of the previous item (issued when ITS pass 1
                // finished, one iteration ago) has long completed, where the order pass 2 -> pass 1 put its ~1.5 k cycles of issue +
                // execution latency on the epilogue's critical path every iteration (trace: acc2 wait 1470 of 7600 cycles)
                pass1(it);
                if (it > 0) pass2(it - 1, pb, pl0, true, b, l0);
            } else {
                if (it > 0) pass2(it - 1, pb, pl0, false, 0, 0);
                {   // residual rows of this tile: in flight under pass 1 and the second MMA
                    const int l = l0 + r;
                    if (l < p.L) {
                        const uint4* src = reinterpret_cast<const uint4*>(p.x_in + ((long)b * p.L + l) * RU_C + hf * 64);
#pragma unroll
                        for (int g = 0; g < 8; ++g) R[g] = src[g];
                    } else {
#pragma unroll
                        for (int g = 0; g < 8; ++g) R[g] = make_uint4(0u, 0u, 0u, 0u);
                    }
                }
                pass1(it);
            }
            pb = b;
            pl0 = l0;
        }
        if (it > 0) pass2(it - 1, pb, pl0, false, 0, 0);
        if (lane == 0) ptx::bulk_wait0();   // every store this thread issued has been performed
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (PAIR) ptx::cluster_sync();   // the peer's shared memory / barriers stay alive until every MMA and remote arrive has landed
    if (warp == 9) {
        ptx::tc_fence_after();
        if (PAIR) ptx::tmem_dealloc_2cta<512>(tmem_base);
        else ptx::tmem_dealloc<512>(tmem_base);
    }
}

}  // namespace

bool vae_resunit128_supported(const VaeResUnit& ru) {
    const VaeConv &c1 = ru.conv1, &c2 = ru.conv2;
    return c1.c_in == RU_C && c1.c_out == RU_C && c2.c_in == RU_C && c2.c_out == RU_C && c1.k == RU_TAPS && c2.k == 1 &&
           c1.w.wtype == W_F16 && c2.w.wtype == W_F16 && c1.w.taps == RU_TAPS && c1.w.kb_per_tap == 2 && c2.w.taps == 1 &&
           c2.w.kb_per_tap == 2 && c1.w.n_tiles == 1 && c2.w.n_tiles == 1 && c1.dilation >= 1 &&
           RU_BM + (RU_TAPS - 1) * c1.dilation <= RU_SLAB_MAX_ROWS && c1.pad == 3 * c1.dilation && c1.bias && c2.bias &&
           ru.snake2.alpha && ru.snake2.inv_beta;
}

cudaError_t launch_vae_resunit128(const __half* xs_in, const __half* x_in, __half* x_out, __half* xs_out, int B, long L,
                                  const VaeResUnit& ru, const VaeSnake& next, cudaStream_t st) {
    if (!vae_resunit128_supported(ru) || !next.alpha || !next.inv_beta || L <= 0 || L > 0x7fffffffL - 4 * RU_BM || (long)B * ((L + RU_BM - 1) / RU_BM) > 0x7fffffffL || xs_in == xs_out)
        return cudaErrorInvalidValue;
    note_launch();
    const double elems = (double)B * (double)L * RU_C;
    ProfScope prof_scope(PROF_GEMM, 2.0 * elems * RU_C * (RU_TAPS + 1), (double)ru.conv1.w.bytes + (double)ru.conv2.w.bytes + elems * 2.0 * (x_out ? 4 : 3), st);
    const int d = ru.conv1.dilation;
    RuParams p;
    p.W1 = static_cast<const uint8_t*>(ru.conv1.w.dev);
    p.W2 = static_cast<const uint8_t*>(ru.conv2.w.dev);
    p.cols[0] = ru.conv1.bias;
    p.cols[1] = ru.snake2.alpha;
    p.cols[2] = ru.snake2.inv_beta;
    p.cols[3] = ru.conv2.bias;
    p.cols[4] = next.alpha;
    p.cols[5] = next.inv_beta;
    p.x_in = x_in;
    p.B = B;
    p.L = (int)L;
    p.m_tiles_per_b = (int)((L + RU_BM - 1) / RU_BM);
    p.dilation = d;
    p.slab_rows = RU_BM + (RU_TAPS - 1) * d;
    p.store_x = x_out ? 1 : 0;
    const uint64_t dims[3] = {(uint64_t)RU_C, (uint64_t)L, (uint64_t)B};
    const uint64_t strides[2] = {(uint64_t)RU_C * 2, (uint64_t)L * RU_C * 2};
    const uint32_t box_a[3] = {64, (uint32_t)p.slab_rows, 1}, box_o[3] = {64, 32, 1};
    CUtensorMap tmA, tmX, tmXS, tmR;
    if (!make_tmap_f16_sw128(&tmA, xs_in, 3, dims, strides, box_a) || !make_tmap_f16_sw128(&tmXS, xs_out, 3, dims, strides, box_o) ||
        !make_tmap_f16_sw128(&tmX, x_out ? x_out : xs_out, 3, dims, strides, box_o) || !make_tmap_f16_sw128(&tmR, x_in, 3, dims, strides, box_o))
        return cudaErrorInvalidValue;
    // one-time per device: shared-memory attributes, and how many CTA pairs can be co-resident (0: clusters unavailable -> single CTA)
    static DeviceOnce setup;
    int max_pairs = 0;
    const cudaError_t ea = setup.run([](int& n) {
        cudaError_t e = cudaFuncSetAttribute(vae_resunit128_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, RuLayout<false>::SMEM);
        if (e != cudaSuccess) return e;
        n = 0;
        if (cudaFuncSetAttribute(vae_resunit128_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, RuLayout<true>::SMEM) == cudaSuccess) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(2 * 148);
            cfg.blockDim = dim3(RU_THREADS);
            cfg.dynamicSmemBytes = RuLayout<true>::SMEM;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = 2;
            at[0].val.clusterDim.y = 1;
            at[0].val.clusterDim.z = 1;
            cfg.attrs = at;
            cfg.numAttrs = 1;
            int c = 0;
            if (cudaOccupancyMaxActiveClusters(&c, vae_resunit128_kernel<true>, &cfg) == cudaSuccess) n = c;
            else cudaGetLastError();
        } else cudaGetLastError();
        return cudaSuccess;
    }, &max_pairs);
    if (ea != cudaSuccess) return ea;
    const long tiles = (long)B * p.m_tiles_per_b;
    const int sms = device_sm_count();
    // ACE_B200_VAE_FUSE=1: single-CTA kernel (A/B and bisecting aid); default: CTA pairs where the device co-schedules enough of them
    const char* fe = std::getenv("ACE_B200_VAE_FUSE");
    const bool pair = !(fe && fe[0] == '1') && max_pairs >= 8 && tiles >= 2;
    p.trace = nullptr;
    p.stall_ns = 0;
    if (const char* es = std::getenv("ACE_B200_RU_STALL_NS")) p.stall_ns = std::max(0, std::min(1000000, std::atoi(es)));
    struct TraceDump {   // debug aid: synchronous, prints the mean timeline of one work item of the leader CTAs
        long long* buf = nullptr; int ctas = 0, step = 1; cudaStream_t st = nullptr; long L = 0; int d = 0, sx = 0, pair = 0;
        ~TraceDump() {
            if (!buf) return;
            cudaStreamSynchronize(st);
            std::vector<long long> h((size_t)ctas * RU_TR_N);
            cudaMemcpy(h.data(), buf, h.size() * sizeof(long long), cudaMemcpyDeviceToHost);
            cudaFree(buf);
            double m[RU_TR_N] = {0};
            int n = 0;
            for (int c = 0; c < ctas; c += step, ++n)
                for (int i = 0; i < RU_TR_N; ++i) m[i] += (double)(h[(size_t)c * RU_TR_N + i] - h[(size_t)c * RU_TR_N]);
            for (double& v : m) v /= n;
            std::fprintf(stderr, "[resunit trace] L=%ld d=%d store_x=%d pair=%d | issuer, cycles from reaching item %d: acc1 free %.0f, slab %.0f, MMA1 issued %.0f, "
                                 "W2 slot %.0f, h ready %.0f, MMA2 issued %.0f\n", L, d, sx, pair, RU_TR_TILE, m[1], m[2], m[3], m[4], m[5], m[6]);
            std::fprintf(stderr, "                 epilogue warp 0: pass2(prev) enters %.0f, acc2 ready %.0f, loaded %.0f, x' out %.0f, snake done %.0f, slice free / residual landed %.0f, "
                                 "xs out %.0f | pass1 enters %.0f, acc1 ready %.0f, h computed %.0f, h written %.0f\n", m[8], m[9], m[10], m[11], m[12], m[13], m[14],
                         m[15], m[16], m[17], m[18]);
        }
    } trace_dump;
    {
        const long items = pair ? (long)B * ((p.m_tiles_per_b + 1) / 2) : tiles;
        const int ctas = pair ? 2 * max_pairs : sms;
        if (std::getenv("ACE_B200_RU_TRACE") && items >= (long)(RU_TR_TILE + 2) * (pair ? max_pairs : sms)) {
            if (cudaMalloc(&trace_dump.buf, (size_t)ctas * RU_TR_N * sizeof(long long)) != cudaSuccess) return cudaErrorMemoryAllocation;
            cudaMemsetAsync(trace_dump.buf, 0, (size_t)ctas * RU_TR_N * sizeof(long long), st);
            p.trace = trace_dump.buf;
            trace_dump.ctas = ctas; trace_dump.step = pair ? 2 : 1; trace_dump.st = st; trace_dump.L = L; trace_dump.d = d; trace_dump.sx = p.store_x; trace_dump.pair = pair;
        }
    }
    if (pair) {
        const long items = (long)B * ((p.m_tiles_per_b + 1) / 2);
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)(2 * (items < max_pairs ? items : max_pairs)));
        cfg.blockDim = dim3(RU_THREADS);
        cfg.dynamicSmemBytes = RuLayout<true>::SMEM;
        cfg.stream = st;
        cudaLaunchAttribute at[2];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 2;
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = 1;
        at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[1].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at;
        cfg.numAttrs = pdl_enabled() ? 2 : 1;
        return cudaLaunchKernelEx(&cfg, vae_resunit128_kernel<true>, tmA, tmX, tmXS, tmR, p);
    }
    return launch_pdl(vae_resunit128_kernel<false>, dim3((unsigned)(tiles < sms ? tiles : sms)), dim3(RU_THREADS), (size_t)RuLayout<false>::SMEM, st, tmA, tmX,
                      tmXS, tmR, p);
}

}  // namespace ace
