// Tap-GEMM for sm_100a (see gemm.h): warp-specialised, persistent, tcgen05 + TMEM + TMA.
//
// CTA = 10 warps:
//   warps 0-3  weight decoders: packed codes (smem, landed by cp.async.bulk) -> fp16 UMMA tile (smem)
//   warps 4-7  epilogue: TMEM -> registers -> fused element-wise -> global
//   warp  8    TMA producer (activation boxes through a tensor map, weight tiles as 1-D bulk copies)
//   warp  9    TMEM allocator + single-thread tcgen05.mma issuer
// Tile = 128 (rows of A) x BN (rows of W) x 64 (K) per pipeline stage; accumulators are double-buffered
// in TMEM (2 x BN columns) so the epilogue of tile i overlaps the MMAs of tile i+1.
#include "gemm.h"
#include "ptx.cuh"
#include "kernels.h"
#include "tmap.h"

#include <cuda.h>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <mutex>
#include <unordered_map>
#include <vector>

namespace ace {

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int A_STAGE_BYTES = BM * BK * 2;     // 16 KB
constexpr int W_SUB_BYTES   = TILE_N * BK * 2; // 16 KB per 128-row sub-tile
constexpr int RAW_MAX_BYTES = 8704;            // largest packed tile (Q8_0)
constexpr int NUM_THREADS   = 320;
constexpr int SLAB_MAX_ROWS = 184;              // 128 + 6 taps x dilation 9 = 182, rounded to whole 8-row swizzle atoms
constexpr int SLAB_PANEL_BYTES = SLAB_MAX_ROWS * 128;  // one 64-channel panel of an activation slab (23 KB, 1 KB aligned)
constexpr int SLAB_BYTES = 2 * SLAB_PANEL_BYTES;       // C <= 128
constexpr int STG_WARP_BYTES = 32 * 32 * 4;    // one epilogue warp's transpose buffer: 32 rows x 32 fp32 columns

struct KParams {
    const uint8_t* W;
    int  wtype, tile_bytes;
    int  N, n_tiles128;          // n_tiles128: number of 128-row weight tiles
    int  kb_per_tap, taps, kb_total;
    int  tap_stride, tap_offset;
    int  B, l_out, m_tiles_per_b;
    int  split_k;
    int  band_m;                 // tile rasterisation: row tiles per band (see tile_coords); >= m_tiles = one band = column-major order
    GemmEpilogue epi;
    int  tma_reduce;             // accumulate epilogue goes out as TMA reduce-add tiles through tmOut
    int  slab_rows;              // SLAB kernel: rows of the activation slab of one tile = 128 + (taps-1) * tap_stride
    int  warm_epi;               // epilogue warps run the epilogue once as a dry run (stores predicated off) while the first tile's MMAs are in flight
    int  epi_variant;            // EPI_STORE fast path chosen on the host (store_variant()); 0 = generic code
    long long* trace;            // debug (ACE_B200_GEMM_TRACE=1): per-CTA clock64 stamps of the pipeline phases, else null
    int  dbg;                    // A/B aid (ACE_B200_GEMM_DBG): bit 1 = CTA-scoped proxy fence in the pair kernel's decoders (the racy one)
    int  qpub;                   // pair kernel, packed tiles: decoders publish their rows with a shared->shared bulk copy (see gemm_tc2_kernel)
};

// trace slots per CTA
enum { TR_START = 0, TR_SETUP = 1, TR_FIRST_FULL = 2, TR_LAST_MMA = 3, TR_ACC_READY = 4, TR_EPI_DONE = 5, TR_END = 6, TR_EPI_LD0 = 7, TR_EPI_CHUNK0 = 8,
       TR_T8_MMA_WAIT = 9, TR_T8_MMA_GO = 10, TR_T8_MMA_DONE = 11, TR_T8_EPI_WAIT = 12, TR_T8_EPI_GO = 13, TR_T8_EPI_DONE = 14, TR_N = 16 };  // T8_*: the CTA's 9th tile (steady state)
__device__ __forceinline__ void trace_stamp(const KParams& p, int slot) {
    if (p.trace) p.trace[(size_t)blockIdx.x * TR_N + slot] = clock64();
}

// Tile rasterisation.  Work item t -> (row tile mt, K split sp, weight tile nt).  Row tiles are grouped in BANDS of band_m: inside a
// band the order is row tile fastest, then split, then weight tile, so the CTAs of a wave share one weight tile and walk the band's
// rows, and the NEXT wave re-reads the SAME activation rows against the next weight tile while they are still in L2.  Without
// bands (one band = all rows) a GEMM whose activations exceed the 126 MB L2 re-streams them from HBM once per weight tile: ncu
// showed 7.5 GB of DRAM reads for the 211 MB of operands of the MLP up-projection at M = 48 000 (profiles/r2_ncu_gemm_config2_k.md).
__device__ __forceinline__ void tile_coords(int t, int m_tiles, int band_m, int nt_sk, int split_k, int& mt, int& sp, int& nt) {
    const int per_band = band_m * nt_sk;
    const int band = t / per_band;
    const int m0 = band * band_m;
    const int bm = min(band_m, m_tiles - m0);   // only the last band can be short
    const int r = t - band * per_band;
    mt = m0 + r % bm;
    const int rest = r / bm;
    sp = rest % split_k;
    nt = rest / split_k;
}

template <int BN, bool QUANT>
struct SmemLayout {
    static constexpr int SUBS       = BN / 128;
    static constexpr int B_STAGE    = SUBS * W_SUB_BYTES;
    static constexpr int RAW_STAGE  = QUANT ? SUBS * RAW_MAX_BYTES : 0;
    static constexpr int STAGE      = A_STAGE_BYTES + B_STAGE;
    static constexpr int EPI_WARPS  = QUANT ? 4 : 8;
    static constexpr int STG_BYTES  = EPI_WARPS * STG_WARP_BYTES;  // epilogue transpose buffers
    static constexpr int COL_BYTES  = 4 * BN * 4;                  // staged bias, gate, Snake alpha, Snake 1/beta of one tile's columns
    static constexpr int BUDGET     = 227 * 1024 - 256 /*barriers*/ - STG_BYTES - COL_BYTES - (QUANT ? 1024 : 0) /*raw-region padding*/;
    static constexpr int STAGES_RAW = BUDGET / (STAGE + RAW_STAGE);
    static constexpr int STAGES     = STAGES_RAW > 6 ? 6 : STAGES_RAW;
    // the epilogue transpose buffers double as the source of TMA reduce-add boxes (128-byte-swizzle image): 1 KB aligned
    static constexpr int RAW_TOTAL  = (STAGES * RAW_STAGE + 1023) / 1024 * 1024;
    static constexpr int BYTES      = STAGES * STAGE + RAW_TOTAL + STG_BYTES + COL_BYTES + 256;
};

// ---------------------------------------------------------------- weight decoders (one thread = one tile row)
__device__ __forceinline__ void st_chunk(uint8_t* btile, int row, int chunk, uint4 v) {
    *reinterpret_cast<uint4*>(btile + row * 128 + ((chunk ^ (row & 7)) << 4)) = v;
}
__device__ __forceinline__ uint32_t h2u(__half2 h) { return *reinterpret_cast<uint32_t*>(&h); }
__device__ __forceinline__ __half2 u2h(uint32_t u) { return *reinterpret_cast<__half2*>(&u); }

// Q8_0 tile: codes [4][128][16] int8, scales [128][2] fp16 at +8192
__device__ __forceinline__ void decode_q8_0(const uint8_t* raw, int row, uint4 (&out)[8]) {
    const __half2 sc2 = *reinterpret_cast<const __half2*>(raw + 8192 + row * 4);
    const __half2 kOff = __floats2half2_rn(1152.f, 1152.f);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const uint4 q = *reinterpret_cast<const uint4*>(raw + j * 2048 + row * 16);
        const __half s = (j < 2) ? __low2half(sc2) : __high2half(sc2);
        const __half2 s2 = __half2half2(s);
        const uint32_t w[4] = {q.x ^ 0x80808080u, q.y ^ 0x80808080u, q.z ^ 0x80808080u, q.w ^ 0x80808080u};
        uint32_t o[8];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            // bytes -> fp16 (1024 + u): 0x64 in the high byte of each half
            const uint32_t lo = __byte_perm(w[i], 0x64646464u, 0x4140);
            const uint32_t hi = __byte_perm(w[i], 0x64646464u, 0x4342);
            o[2 * i]     = h2u(__hmul2(__hsub2(u2h(lo), kOff), s2));
            o[2 * i + 1] = h2u(__hmul2(__hsub2(u2h(hi), kOff), s2));
        }
        out[2 * j] = make_uint4(o[0], o[1], o[2], o[3]);
        out[2 * j + 1] = make_uint4(o[4], o[5], o[6], o[7]);
    }
}

// Q4_K tile: nibble words [2][128][16], (d*sc, dmin*m) fp16 pairs [128][2] at +4096
__device__ __forceinline__ void decode_q4_k(const uint8_t* raw, int row, uint4 (&out)[8]) {
    const uint2 sm = *reinterpret_cast<const uint2*>(raw + 4096 + row * 8);
    const __half2 k1024 = __floats2half2_rn(1024.f, 1024.f);
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        const uint4 q = *reinterpret_cast<const uint4*>(raw + j * 2048 + row * 16);
        const __half2 pr = u2h(j == 0 ? sm.x : sm.y);
        const __half2 ds = __half2half2(__low2half(pr));
        const __half2 dm = __hneg2(__half2half2(__high2half(pr)));
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int wi = 0; wi < 4; ++wi) {
            uint32_t o[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const uint32_t t = ((w[wi] >> (4 * i)) & 0x000F000Fu) | 0x64006400u;
                o[i] = h2u(__hfma2(__hsub2(u2h(t), k1024), ds, dm));
            }
            out[4 * j + wi] = make_uint4(o[0], o[1], o[2], o[3]);
        }
    }
}

// Q6_K tile: low nibble words [2][128][16], 2-bit plane [128][16] at +4096, fp16 d*sc [128][4] at +6144
__device__ __forceinline__ void decode_q6_k(const uint8_t* raw, int row, uint4 (&out)[8]) {
    const uint4 hp = *reinterpret_cast<const uint4*>(raw + 4096 + row * 16);
    const uint2 sc = *reinterpret_cast<const uint2*>(raw + 6144 + row * 8);
    const uint32_t P[4] = {hp.x, hp.y, hp.z, hp.w};
    const __half2 k1056 = __floats2half2_rn(1056.f, 1056.f);
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        const uint4 q = *reinterpret_cast<const uint4*>(raw + j * 2048 + row * 16);
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int wi = 0; wi < 4; ++wi) {
            const int g = 4 * j + wi;
            const uint32_t Hs = (g & 1) ? (P[g >> 1] >> 4) : (P[g >> 1] << 4);
            const uint32_t scw = (g >> 1) < 2 ? sc.x : sc.y;
            const __half sh = ((g >> 1) & 1) ? __high2half(u2h(scw)) : __low2half(u2h(scw));
            const __half2 s2 = __half2half2(sh);
            uint32_t o[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const uint32_t t = ((w[wi] >> (4 * i)) & 0x000F000Fu) | ((Hs >> (2 * i)) & 0x00300030u) | 0x64006400u;
                o[i] = h2u(__hmul2(__hsub2(u2h(t), k1056), s2));
            }
            out[g] = make_uint4(o[0], o[1], o[2], o[3]);
        }
    }
}

// one row of a packed tile -> its 64 fp16 weights as eight 16-byte chunks (registers)
__device__ __forceinline__ void decode_row(int wtype, const uint8_t* raw, int row, uint4 (&out)[8]) {
    if (wtype == W_Q8_0) decode_q8_0(raw, row, out);
    else if (wtype == W_Q4_K) decode_q4_k(raw, row, out);
    else decode_q6_k(raw, row, out);
}
// ... and into the 128-byte-swizzle K-major image of the tile
__device__ __forceinline__ void store_row(uint8_t* btile, int row, const uint4 (&out)[8]) {
#pragma unroll
    for (int c = 0; c < 8; ++c) st_chunk(btile, row, c, out[c]);
}
__device__ __forceinline__ void decode_tile(int wtype, const uint8_t* raw, uint8_t* btile, int row) {
    uint4 out[8];
    decode_row(wtype, raw, row, out);
    store_row(btile, row, out);
}

// ---------------------------------------------------------------- fused epilogue for 32 consecutive columns of one row
// Stores that a run-time flag can switch off WITHOUT changing the instruction stream: the epilogue warm-up pass (`warm_epi`)
// executes the real code with `live` = false, so the code is fetched and the loads are issued, but nothing is written.
// No memory clobber: the values stored depend on every load they must follow, and nothing re-reads these addresses.
template <bool W>
__device__ __forceinline__ void st_v4_f32_if(bool live, float* ptr, float4 v) {
    if (!W) {
        *reinterpret_cast<float4*>(ptr) = v;
        return;
    }
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %0, 0;\n\t@p st.global.v4.f32 [%1], {%2, %3, %4, %5};\n\t}" ::"r"((int)live), "l"(ptr), "f"(v.x),
                 "f"(v.y), "f"(v.z), "f"(v.w));
}
template <bool W>
__device__ __forceinline__ void st_v2_b32_if(bool live, void* ptr, uint32_t x, uint32_t y) {
    if (!W) {
        *reinterpret_cast<uint2*>(ptr) = make_uint2(x, y);
        return;
    }
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %0, 0;\n\t@p st.global.v2.b32 [%1], {%2, %3};\n\t}" ::"r"((int)live), "l"(ptr), "r"(x), "r"(y));
}
template <bool W>
__device__ __forceinline__ void st_v4_b32_if(bool live, void* ptr, uint4 v) {
    if (!W) {
        *reinterpret_cast<uint4*>(ptr) = v;
        return;
    }
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %0, 0;\n\t@p st.global.v4.b32 [%1], {%2, %3, %4, %5};\n\t}" ::"r"((int)live), "l"(ptr), "r"(v.x),
                 "r"(v.y), "r"(v.z), "r"(v.w));
}

// Snake: v + sin^2(a v) / b through the MUFU sine and its own range reduction: abs phase error ~|a v| * 1e-7 rad, far below
// the fp16 rounding of the activation it produces.  5 instructions per element (a two-term Cody-Waite reduction in front of
// the MUFU cost 10 and made the VAE epilogues issue-bound).
__device__ __forceinline__ float snake_f(float v, float a, float inv_b) {
    const float s = __sinf(a * v);
    return fmaf(s * s, inv_b, v);
}

template <bool W>
__device__ __forceinline__ void epilogue_row32(const KParams& p, const float (&acc)[32], int b, int l, int n0,
                                               bool add_bias, bool dry) {
    const GemmEpilogue& e = p.epi;
    if (l >= p.l_out && !(W && dry && e.mode == EPI_SWIGLU)) return;
    if (e.mode == EPI_SWIGLU) {
        // columns (2j, 2j+1) = (gate_j, up_j)
        const int nout = p.N >> 1;
        const int j0 = n0 >> 1;
        __half* o = e.out_f16 + (long)b * e.out_bstride + (long)l * e.ld_out + j0;
        __align__(16) __half h[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const float g = acc[2 * i], u = acc[2 * i + 1];
            h[i] = __float2half_rn(__fdividef(g, 1.0f + __expf(-g)) * u);  // fast divide: the result is rounded to fp16
        }
        const bool live = !(W && dry) && l < p.l_out;
        if (j0 + 16 <= nout) {
            st_v4_b32_if<W>(live, o, *reinterpret_cast<const uint4*>(h));
            st_v4_b32_if<W>(live, o + 8, *reinterpret_cast<const uint4*>(h + 8));
        } else if (live) {
            for (int i = 0; i < 16 && j0 + i < nout; ++i) o[i] = h[i];
        }
        return;
    }
    if (W && dry) return;
    const long flat0 = (long)l * e.ld_out + n0 + e.out_shift;  // position inside the sample
    const long base = (long)b * e.out_bstride + flat0;
    const int ncols = min(32, p.N - n0);
    if (ncols <= 0) return;
    const bool interior = (ncols == 32) && flat0 >= 0 && flat0 + 32 <= e.out_valid;
    float v[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = acc[i];
    if (e.bias && add_bias) {
#pragma unroll
        for (int i = 0; i < 32; ++i) if (i < ncols) v[i] += __ldg(e.bias + n0 + i);
    }
    if (e.gate) {
        const float* g = e.gate + (long)b * e.gate_bstride + n0;
#pragma unroll
        for (int i = 0; i < 32; ++i) if (i < ncols) v[i] *= __ldg(g + i);
    }
    if (e.atomic) {
        if (interior && ((base & 3) == 0)) {  // vector reduction: 8 x red.v4.f32 instead of 32 scalar atomics
            float* o = e.out_f32 + base;
#pragma unroll
            for (int i = 0; i < 8; ++i)
                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(o + 4 * i), "f"(v[4 * i]), "f"(v[4 * i + 1]),
                             "f"(v[4 * i + 2]), "f"(v[4 * i + 3])
                             : "memory");
            return;
        }
        for (int i = 0; i < ncols; ++i) {
            const long f = flat0 + i;
            if (f >= 0 && f < e.out_valid) atomicAdd(e.out_f32 + base + i, v[i]);
        }
        return;
    }
    if (interior && ((base & 3) == 0)) {
        if (e.resid) {
            const float4* r = reinterpret_cast<const float4*>(e.resid + base);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float4 t = r[i];
                v[4 * i] += t.x; v[4 * i + 1] += t.y; v[4 * i + 2] += t.z; v[4 * i + 3] += t.w;
            }
        }
        if (e.resid16) {
            const uint2* r = reinterpret_cast<const uint2*>(e.resid16 + base);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const uint2 t = r[i];
                const float2 a = __half22float2(u2h(t.x)), b2 = __half22float2(u2h(t.y));
                v[4 * i] += a.x; v[4 * i + 1] += a.y; v[4 * i + 2] += b2.x; v[4 * i + 3] += b2.y;
            }
        }
        if (e.out_x16) {
            uint2* o = reinterpret_cast<uint2*>(e.out_x16 + base);
#pragma unroll
            for (int i = 0; i < 8; ++i)
                o[i] = make_uint2(h2u(__floats2half2_rn(v[4 * i], v[4 * i + 1])), h2u(__floats2half2_rn(v[4 * i + 2], v[4 * i + 3])));
        }
        if (e.out_f32) {
            float4* o = reinterpret_cast<float4*>(e.out_f32 + base);
#pragma unroll
            for (int i = 0; i < 8; ++i) o[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        }
        if (e.out_f16) {
            if (e.snake_alpha) {
                const int c0 = e.snake_period ? (n0 % e.snake_period) : n0;  // period is a multiple of 32 or N<=period
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = snake_f(v[i], __ldg(e.snake_alpha + c0 + i), __ldg(e.snake_invbeta + c0 + i));
            } else if (e.silu_f16) {
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = __fdividef(v[i], 1.0f + __expf(-v[i]));
            }
            __align__(16) __half h[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) h[i] = __float2half_rn(v[i]);
            if ((base & 7) == 0) {
                uint4* o = reinterpret_cast<uint4*>(e.out_f16 + base);
#pragma unroll
                for (int i = 0; i < 4; ++i) o[i] = reinterpret_cast<const uint4*>(h)[i];
            } else {
                uint2* o = reinterpret_cast<uint2*>(e.out_f16 + base);
#pragma unroll
                for (int i = 0; i < 8; ++i) o[i] = reinterpret_cast<const uint2*>(h)[i];
            }
        }
        return;
    }
    // edge path: element-wise with bounds checks
    for (int i = 0; i < ncols; ++i) {
        const long f = flat0 + i;
        if (f < 0 || f >= e.out_valid) continue;
        float x = v[i];
        if (e.resid) x += e.resid[base + i];
        if (e.resid16) x += __half2float(e.resid16[base + i]);
        if (e.out_x16) e.out_x16[base + i] = __float2half_rn(x);
        if (e.out_f32) e.out_f32[base + i] = x;
        if (e.out_f16) {
            if (e.snake_alpha) {
                const int c = e.snake_period ? ((n0 + i) % e.snake_period) : (n0 + i);
                x = snake_f(x, __ldg(e.snake_alpha + c), __ldg(e.snake_invbeta + c));
            } else if (e.silu_f16) {
                x = __fdividef(x, 1.0f + __expf(-x));
            }
            e.out_f16[base + i] = __float2half_rn(x);
        }
    }
}

// Specialised phase 2 of the staged EPI_STORE chunk for a chunk that lies entirely inside the output (no per-row predicates,
// no run-time feature tests in the loop).  ncu on the last-level VAE convolutions showed the generic code at ~1200 warp
// instructions per 32x32 chunk — issue-bound, 8.5 k cycles of epilogue per 128x128 tile against 3.6 k of MMAs
// (profiles/r1_vae_conv_ncu.md); this form needs ~350.
enum { SV_GENERIC = 0, SV_F16 = 1, SV_F16_SNAKE = 2, SV_F16_SILU = 3, SV_F32 = 4, SV_RES_F32_F16SNAKE = 5, SV_RES_F16SNAKE = 6, SV_F32_F16SNAKE = 7,
       SV_RES16_X16_F16SNAKE = 8, SV_RES16_F16SNAKE = 9, SV_X16_F16SNAKE = 10 };  // 8..10: 16-bit residual stream (resid16 / out_x16)
// RESID: 0 none, 1 fp32 `resid`, 2 fp16 `resid16`.  OUTX: 0 none, 1 fp32 `out_f32`, 2 fp16 `out_x16`.
template <bool W, int ACT, int RESID, int OUTX, bool OUT16>
__device__ __forceinline__ void staged_phase2_fast(const GemmEpilogue& e, const float4* s4, long off0, long step, int lane, int n0, bool live,
                                                   const float* s_sa, const float* s_sb) {
    const int seg = lane & 7, r0 = lane >> 3;
    float4 res[8];
    if (RESID == 1) {
#pragma unroll
        for (int k = 0; k < 8; ++k) res[k] = __ldcg(reinterpret_cast<const float4*>(e.resid + off0 + k * step));
    } else if (RESID == 2) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const uint2 t = __ldcg(reinterpret_cast<const uint2*>(e.resid16 + off0 + k * step));
            const float2 a = __half22float2(u2h(t.x)), b2 = __half22float2(u2h(t.y));
            res[k] = make_float4(a.x, a.y, b2.x, b2.y);
        }
    }
    float4 sa = make_float4(0.f, 0.f, 0.f, 0.f), sb = sa;
    if (ACT == 1) {
        if (s_sa) {  // staged in shared memory with the tile's bias (single-CTA kernel)
            sa = *reinterpret_cast<const float4*>(s_sa + 4 * seg);
            sb = *reinterpret_cast<const float4*>(s_sb + 4 * seg);
        } else {
            const int c0 = (e.snake_period ? (n0 % e.snake_period) : n0) + 4 * seg;
            sa = __ldg(reinterpret_cast<const float4*>(e.snake_alpha + c0));
            sb = __ldg(reinterpret_cast<const float4*>(e.snake_invbeta + c0));
        }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int rr = 4 * k + r0;
        float4 t = s4[rr * 8 + (seg ^ (rr & 7))];
        if (RESID) { t.x += res[k].x; t.y += res[k].y; t.z += res[k].z; t.w += res[k].w; }
        if (OUTX == 1) st_v4_f32_if<W>(live, e.out_f32 + off0 + k * step, t);
        if (OUTX == 2) st_v2_b32_if<W>(live, e.out_x16 + off0 + k * step, h2u(__floats2half2_rn(t.x, t.y)), h2u(__floats2half2_rn(t.z, t.w)));
        if (OUT16) {
            if (ACT == 1) {
                t.x = snake_f(t.x, sa.x, sb.x); t.y = snake_f(t.y, sa.y, sb.y);
                t.z = snake_f(t.z, sa.z, sb.z); t.w = snake_f(t.w, sa.w, sb.w);
            } else if (ACT == 2) {
                t.x = __fdividef(t.x, 1.0f + __expf(-t.x)); t.y = __fdividef(t.y, 1.0f + __expf(-t.y));
                t.z = __fdividef(t.z, 1.0f + __expf(-t.z)); t.w = __fdividef(t.w, 1.0f + __expf(-t.w));
            }
            const __half2 h0 = __floats2half2_rn(t.x, t.y), h1 = __floats2half2_rn(t.z, t.w);
            st_v2_b32_if<W>(live, e.out_f16 + off0 + k * step, h2u(h0), h2u(h1));
        }
    }
}

// EPI_STORE for a full 32-column chunk, through the warp's shared-memory transpose buffer.
// Phase 1 (thread = accumulator row, as tcgen05.ld delivers it) applies the per-column terms (bias, AdaLN gate) and parks
// the row in smem; phase 2 re-reads it so that 8 lanes cover one row's 128 contiguous bytes and one warp instruction
// touches 4 full 128-byte lines — instead of 32 rows x 16 bytes, which made every epilogue cost ~12 k cycles per 128x256
// tile (profiles/r1_gemm_trace_tr1.log) because each 16-byte piece was its own L2 transaction.  The residual read, the fp32
// store / red.add and the fp16 (Snake / SiLU) store all happen in phase 2.  16-byte pieces are XOR-swizzled by the row so
// both phases are bank-conflict free.  Returns false (nothing done) when the chunk does not qualify: caller falls back.
template <bool W>
__device__ __forceinline__ bool epilogue_store_staged(const KParams& p, const CUtensorMap* tmOut, const float (&acc)[32], int b, int l_base,
                                                      int lane, int n0, const float* s_bias, const float* s_gate, float* stg, bool dry,
                                                      const float* s_sa = nullptr, const float* s_sb = nullptr) {
    const GemmEpilogue& e = p.epi;
    if (n0 + 32 > p.N) return false;                                     // warp-uniform
    const long col0 = (long)n0 + e.out_shift;
    if (((e.ld_out | col0 | e.out_bstride) & 3) != 0) return false;      // 16-byte alignment of every row segment
    if (p.tma_reduce) {  // the previous chunk's reduce-add has finished READING this buffer (its global side may still run)
        if (lane == 0) ptx::bulk_wait_read0();
        __syncwarp();
    }
    // per-column terms of this chunk, staged in smem once per tile (s_bias / s_gate point at this chunk's 32 entries)
    float4* s4 = reinterpret_cast<float4*>(stg);
    if (e.gate) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float4 bz = reinterpret_cast<const float4*>(s_bias)[i], gz = reinterpret_cast<const float4*>(s_gate)[i];
            s4[lane * 8 + (i ^ (lane & 7))] = make_float4((acc[4 * i] + bz.x) * gz.x, (acc[4 * i + 1] + bz.y) * gz.y,
                                                          (acc[4 * i + 2] + bz.z) * gz.z, (acc[4 * i + 3] + bz.w) * gz.w);
        }
    } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float4 bz = reinterpret_cast<const float4*>(s_bias)[i];
            s4[lane * 8 + (i ^ (lane & 7))] = make_float4(acc[4 * i] + bz.x, acc[4 * i + 1] + bz.y, acc[4 * i + 2] + bz.z, acc[4 * i + 3] + bz.w);
        }
    }
    if (p.tma_reduce) {
        // x += tile: one TMA reduce-add of the 32 x 32 fp32 box per chunk (the staging layout IS the 128-byte-swizzle box
        // image); rows >= l_out and columns >= N are clipped by the tensor map.  Replaces 1024 per-element red ops whose
        // per-SM issue rate (~2.2 k cycles per chunk) was the exposed part of the small-M GEMMs.
        ptx::fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0 && !(W && dry)) {
            ptx::tma_reduce_add_3d(tmOut, stg, n0, l_base, b);
            ptx::bulk_commit();
        }
        return true;
    }
    __syncwarp();
    if (p.epi_variant != SV_GENERIC) {
        // whole chunk inside the output?  (rows ascend with l, so the first and the last row bound the flat range)
        const long first = (long)l_base * e.ld_out + col0, last = first + 31 * e.ld_out;
        if (l_base + 32 <= p.l_out && first >= 0 && last + 32 <= e.out_valid) {
            const long off0 = (long)b * e.out_bstride + first + (long)(lane >> 3) * e.ld_out + 4 * (lane & 7);
            const long step = 4 * e.ld_out;
            const bool live = !(W && dry);
            switch (p.epi_variant) {
                case SV_F16:              staged_phase2_fast<W, 0, 0, 0, true>(e, s4, off0, step, lane, n0, live, s_sa, s_sb); break;
                case SV_F16_SNAKE:        staged_phase2_fast<W, 1, 0, 0, true>(e, s4, off0, step, lane, n0, live, s_sa, s_sb); break;
                case SV_F16_SILU:         staged_phase2_fast<W, 2, 0, 0, true>(e, s4, off0, step, lane, n0, live, s_sa, s_sb); break;
                case SV_F32:              staged_phase2_fast<W, 0, 0, 1, false>(e, s4, off0, step, lane, n0, live, s_sa, s_sb); break;
                case SV_RES_F32_F16SNAKE: staged_phase2_fast<W, 1, 1, 1, true>(e, s4, off0, step, lane, n0, live, s_sa, s_sb); break;
                case SV_RES_F16SNAKE:     staged_phase2_fast<W, 1, 1, 0, true>(e, s4, off0, step, lane, n0, live, s_sa, s_sb); break;
                case SV_RES16_X16_F16SNAKE: staged_phase2_fast<W, 1, 2, 2, true>(e, s4, off0, step, lane, n0, live, s_sa, s_sb); break;
                case SV_RES16_F16SNAKE:   staged_phase2_fast<W, 1, 2, 0, true>(e, s4, off0, step, lane, n0, live, s_sa, s_sb); break;
                case SV_X16_F16SNAKE:     staged_phase2_fast<W, 1, 0, 2, true>(e, s4, off0, step, lane, n0, live, s_sa, s_sb); break;
                default:                  staged_phase2_fast<W, 1, 0, 1, true>(e, s4, off0, step, lane, n0, live, s_sa, s_sb); break;
            }
            __syncwarp();  // buffer is reused by the next chunk
            return true;
        }
    }
    if (W && dry) return true;  // the generic tail is not worth warming
    const int seg = lane & 7;                 // this lane's 4 columns: n0 + 4 seg ..
    long off[8];
    bool ok[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int l = l_base + 4 * k + (lane >> 3);
        const long flat0 = (long)l * e.ld_out + col0;   // chunk start inside the sample
        ok[k] = l < p.l_out && flat0 >= 0 && flat0 + 32 <= e.out_valid;
        off[k] = (long)b * e.out_bstride + flat0 + 4 * seg;
    }
    if (e.atomic) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int rr = 4 * k + (lane >> 3);
            const float4 t = s4[rr * 8 + (seg ^ (rr & 7))];
            if (ok[k])
                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(e.out_f32 + off[k]), "f"(t.x), "f"(t.y), "f"(t.z), "f"(t.w) : "memory");
        }
        __syncwarp();
        return true;
    }
    float4 res[8];
    if (e.resid) {  // all eight row segments in flight before the first use
#pragma unroll
        for (int k = 0; k < 8; ++k) res[k] = ok[k] ? __ldcg(reinterpret_cast<const float4*>(e.resid + off[k])) : make_float4(0.f, 0.f, 0.f, 0.f);
    } else if (e.resid16) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            res[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (ok[k]) {
                const uint2 t = __ldcg(reinterpret_cast<const uint2*>(e.resid16 + off[k]));
                const float2 a = __half22float2(u2h(t.x)), b2 = __half22float2(u2h(t.y));
                res[k] = make_float4(a.x, a.y, b2.x, b2.y);
            }
        }
    }
    float4 sa = make_float4(0.f, 0.f, 0.f, 0.f), sb = sa;
    if (e.out_f16 && e.snake_alpha) {
        if (s_sa) {
            sa = *reinterpret_cast<const float4*>(s_sa + 4 * seg);
            sb = *reinterpret_cast<const float4*>(s_sb + 4 * seg);
        } else {
            const int c0 = (e.snake_period ? (n0 % e.snake_period) : n0) + 4 * seg;
            sa = __ldg(reinterpret_cast<const float4*>(e.snake_alpha + c0));
            sb = __ldg(reinterpret_cast<const float4*>(e.snake_invbeta + c0));
        }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int rr = 4 * k + (lane >> 3);
        float4 t = s4[rr * 8 + (seg ^ (rr & 7))];
        if (!ok[k]) continue;
        if (e.resid || e.resid16) { t.x += res[k].x; t.y += res[k].y; t.z += res[k].z; t.w += res[k].w; }
        if (e.out_f32) *reinterpret_cast<float4*>(e.out_f32 + off[k]) = t;
        if (e.out_x16) *reinterpret_cast<uint2*>(e.out_x16 + off[k]) = make_uint2(h2u(__floats2half2_rn(t.x, t.y)), h2u(__floats2half2_rn(t.z, t.w)));
        if (e.out_f16) {
            if (e.snake_alpha) {
                t.x = snake_f(t.x, sa.x, sb.x); t.y = snake_f(t.y, sa.y, sb.y);
                t.z = snake_f(t.z, sa.z, sb.z); t.w = snake_f(t.w, sa.w, sb.w);
            } else if (e.silu_f16) {
                t.x = __fdividef(t.x, 1.0f + __expf(-t.x)); t.y = __fdividef(t.y, 1.0f + __expf(-t.y));
                t.z = __fdividef(t.z, 1.0f + __expf(-t.z)); t.w = __fdividef(t.w, 1.0f + __expf(-t.w));
            }
            const __half2 h0 = __floats2half2_rn(t.x, t.y), h1 = __floats2half2_rn(t.z, t.w);
            *reinterpret_cast<uint2*>(e.out_f16 + off[k]) = make_uint2(h2u(h0), h2u(h1));
        }
    }
    __syncwarp();  // buffer is reused by the next chunk
    return true;
}

// Column vectors of one tile for the staged epilogue: bias (0 when absent / not this split's job) and AdaLN gate (1 when
// absent) of the tile's BN columns, fetched by the epilogue warps while the tile's MMAs are still running.
template <int BN, int NTHREADS>
__device__ __forceinline__ void stage_column_terms(const KParams& p, int b, int n_base, bool add_bias, int tid, float* s_bias, float* s_gate,
                                                   float* s_snake = nullptr) {
    const GemmEpilogue& e = p.epi;
    for (int j = tid; j < BN; j += NTHREADS) {
        const int n = n_base + j;
        const bool in = n < p.N;
        s_bias[j] = (in && e.bias && add_bias) ? __ldg(e.bias + n) : 0.f;
        s_gate[j] = (in && e.gate) ? __ldg(e.gate + (long)b * e.gate_bstride + n) : 1.f;
        if (s_snake && e.snake_alpha) {  // per-channel Snake constants: ncu had the phase-2 loads of these at 9 % of all stall samples
            const int ch = e.snake_period ? (n % e.snake_period) : n;
            s_snake[j] = in ? __ldg(e.snake_alpha + ch) : 0.f;
            s_snake[BN + j] = in ? __ldg(e.snake_invbeta + ch) : 0.f;
        }
    }
}
// EPI_HEADNORM: q/k norm weights of the tile's heads (128 floats per head, zeros for V heads) into smem, under the MMAs.
template <int BN, int NTHREADS>
__device__ __forceinline__ void stage_head_weights(const KParams& p, int n_base, int tid, float* s_w) {
    const GemmEpilogue& e = p.epi;
    for (int jx = tid; jx < BN; jx += NTHREADS) {
        const int n0 = n_base + (jx & ~127);
        const float* w = n0 < e.hn_na ? e.hn_wa : (n0 < e.hn_na + e.hn_nb ? e.hn_wb : nullptr);
        s_w[jx] = w ? __ldg(w + (jx & 127)) : 0.f;
    }
}
__device__ __forceinline__ void epi_bar_sync(int nthreads) {  // named barrier 1: the epilogue warps only
    asm volatile("bar.sync 1, %0;" ::"r"(nthreads) : "memory");
}

// EPI_HEADNORM: one head (128 accumulator columns) of the warp's 32 rows.  Pass 1 (thread = accumulator row) reads the four
// 32-column chunks from TMEM for the sum of squares.  Pass 2 re-reads them as rotate-half pairs (chunk c with chunk c+2),
// applies inv-rms x norm weight, and parks 16 + 16 columns per row in the warp's transpose buffer; phase 2 then gives every
// row to 4 lanes (8 rows per step), so the RoPE table is read and the fp16 head written as 128 / 32 contiguous bytes per row:
// 8 lines per warp instruction instead of 32.  The thread-per-row form was L1-request bound: 64 loads x 32 lines per head and
// warp = ~17 k cycles of epilogue behind a 19 k cycle main loop (profiles/r1_trace_dit_hn.log).
// `c_begin / c_step` let two warp groups split one head by chunk pair (128-column tiles have a single head).
template <bool W>
__device__ __forceinline__ void epilogue_head128(const KParams& p, uint32_t taddr, int b, bool live, int l_base, int lane, int n0,
                                                 float* stg, const float* s_w, int c_begin, int c_step) {
    const GemmEpilogue& e = p.epi;
    const float* w = (n0 < e.hn_na + e.hn_nb) ? s_w : nullptr;  // norm weights of this head, staged in smem by stage_head_weights
    uint32_t r0[32], r1[32];
    float inv = 1.f;
    if (w) {
        float ss = 0.f;
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
            ptx::tmem_ld32(taddr + c * 32, r0);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const float v = __uint_as_float(r0[i]);
                ss = fmaf(v, v, ss);
            }
        }
        inv = rsqrtf(ss * (1.0f / 128.0f) + e.hn_eps);
    }
    const bool rope = w && e.rope;
    float4* s4 = reinterpret_cast<float4*>(stg);
    const int j = lane & 3, rq = lane >> 2;
#pragma unroll 1
    for (int c = c_begin; c < 2; c += c_step) {
        ptx::tmem_ld32(taddr + c * 32, r0);
        ptx::tmem_ld32(taddr + (c + 2) * 32, r1);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            // this lane's RoPE entries of the four row groups, requested before the staging work so the L2 round trip
            // (L1 is nearly all shared memory here) overlaps it
            float4 cs[4][2];
            if (rope) {
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int l = min(l_base + 8 * k + rq, p.l_out - 1);
                    const float4* rp = reinterpret_cast<const float4*>(e.rope + (long)l * 64 + c * 32 + h * 16 + 4 * j);
                    cs[k][0] = __ldg(rp);
                    cs[k][1] = __ldg(rp + 1);
                }
            }
#pragma unroll
            for (int i4 = 0; i4 < 4; ++i4) {
                const int col = h * 16 + 4 * i4;
                float4 av = make_float4(__uint_as_float(r0[col]), __uint_as_float(r0[col + 1]), __uint_as_float(r0[col + 2]), __uint_as_float(r0[col + 3]));
                float4 bv = make_float4(__uint_as_float(r1[col]), __uint_as_float(r1[col + 1]), __uint_as_float(r1[col + 2]), __uint_as_float(r1[col + 3]));
                if (w) {
                    const float4 wa = *reinterpret_cast<const float4*>(w + c * 32 + col);
                    const float4 wb = *reinterpret_cast<const float4*>(w + 64 + c * 32 + col);
                    av.x *= inv * wa.x; av.y *= inv * wa.y; av.z *= inv * wa.z; av.w *= inv * wa.w;
                    bv.x *= inv * wb.x; bv.y *= inv * wb.y; bv.z *= inv * wb.z; bv.w *= inv * wb.w;
                }
                s4[lane * 8 + (i4 ^ (lane & 7))] = av;
                s4[lane * 8 + ((4 + i4) ^ (lane & 7))] = bv;
            }
            __syncwarp();
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int rr = 8 * k + rq;
                const int l = l_base + rr;
                float4 av = s4[rr * 8 + (j ^ (rr & 7))], bv = s4[rr * 8 + ((4 + j) ^ (rr & 7))];
                if (l < p.l_out) {
                    const int col = c * 32 + h * 16 + 4 * j;
                    if (rope) {
                        const float4 cs01 = cs[k][0], cs23 = cs[k][1];
                        float t;
                        t = av.x * cs01.x - bv.x * cs01.y; bv.x = av.x * cs01.y + bv.x * cs01.x; av.x = t;
                        t = av.y * cs01.z - bv.y * cs01.w; bv.y = av.y * cs01.w + bv.y * cs01.z; av.y = t;
                        t = av.z * cs23.x - bv.z * cs23.y; bv.z = av.z * cs23.y + bv.z * cs23.x; av.z = t;
                        t = av.w * cs23.z - bv.w * cs23.w; bv.w = av.w * cs23.w + bv.w * cs23.z; av.w = t;
                    }
                    __half* o = e.out_f16 + (long)b * e.out_bstride + (long)l * e.ld_out + n0 + col;
                    st_v2_b32_if<W>(live, o, h2u(__floats2half2_rn(av.x, av.y)), h2u(__floats2half2_rn(av.z, av.w)));
                    st_v2_b32_if<W>(live, o + 64, h2u(__floats2half2_rn(bv.x, bv.y)), h2u(__floats2half2_rn(bv.z, bv.w)));
                }
            }
            __syncwarp();
        }
    }
}

}  // namespace

namespace {

// ---------------------------------------------------------------- the tensor-core kernel
template <int BN, bool QUANT, bool WARM, bool SLAB = false>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmOut, const KParams p) {
    using SL = SmemLayout<BN, QUANT>;
    constexpr int STAGES = SL::STAGES;
    constexpr int SUBS = SL::SUBS;
    constexpr int TMEM_COLS = 2 * BN;  // two accumulators

    if (threadIdx.x == 0) trace_stamp(p, TR_START);
    extern __shared__ __align__(1024) uint8_t smem[];
    if (threadIdx.x == 0 && (ptx::smem_u32(smem) & 1023u) != 0) __trap();  // the 128-byte swizzle needs 1 KB aligned tiles
    // SLAB (k7 convolutions, C <= 128): instead of one 128-row activation box per k-block, the 128 + (taps-1)*dilation rows a
    // tile needs are loaded ONCE as 64-channel panels (double-buffered across tiles) and tap t reads them at row offset
    // t*dilation: a K-major SWIZZLE_128B operand may start at any 128-byte row of a TMA-written panel because the swizzle is
    // a function of the shared-memory address (scripts/probes/umma_rowshift_probe.cu: exact for every shift, base_offset 0).
    uint8_t* sA = smem;                                  // [STAGES][16 KB]   |  SLAB: [2][SLAB_BYTES]
    uint8_t* sB = sA + (SLAB ? 2 * SLAB_BYTES : STAGES * A_STAGE_BYTES);  // [STAGES][SUBS*16 KB]
    uint8_t* sR = sB + STAGES * SL::B_STAGE;             // [STAGES][SUBS*RAW_MAX] (QUANT only)
    float* sStg = reinterpret_cast<float*>(sR + SL::RAW_TOTAL);  // [EPI_WARPS][32 x 32 fp32], 1 KB aligned
    float* sCol = sStg + SL::STG_BYTES / 4;  // [4][BN]: bias, gate, Snake alpha, Snake 1/beta of the current tile's columns
    uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(sCol) + SL::COL_BYTES);
    uint64_t* full = bars;                  // [STAGES] operands of stage s are in smem
    uint64_t* empty = bars + STAGES;        // [STAGES] MMAs reading stage s have completed
    uint64_t* raw_full = bars + 2 * STAGES; // [STAGES] packed codes landed
    uint64_t* raw_empty = bars + 3 * STAGES;// [STAGES] decoders done with the packed codes
    uint64_t* acc_full = bars + 4 * STAGES; // [2]
    uint64_t* acc_empty = acc_full + 2;     // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    const int m_tiles = p.B * p.m_tiles_per_b;
    const int n_tiles = (p.n_tiles128 + SUBS - 1) / SUBS;
    const int total_tiles = n_tiles * p.split_k * m_tiles;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            ptx::mbar_init(&full[s], QUANT ? 1 + 4 : 1);
            ptx::mbar_init(&empty[s], 1);
            ptx::mbar_init(&raw_full[s], 1);
            ptx::mbar_init(&raw_empty[s], 4);
        }
        if (SLAB) {  // raw_full[0..1] / raw_empty[0..1] serve as slab_full / slab_empty
            ptx::mbar_init(&raw_empty[0], 1);
            ptx::mbar_init(&raw_empty[1], 1);
        }
        for (int a = 0; a < 2; ++a) {
            ptx::mbar_init(&acc_full[a], 1);
            ptx::mbar_init(&acc_empty[a], QUANT ? 4 : 8);
        }
        ptx::fence_mbar_init();
    }
    if (warp == 9) {
        ptx::tmem_alloc<TMEM_COLS>(tmem_slot);
    }
    if (warp == 8 && lane == 0) ptx::prefetch_tmap(&tmA);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    ptx::pdl_wait();    // everything above overlapped the previous kernel's tail
    ptx::pdl_launch();
    if (threadIdx.x == 0) trace_stamp(p, TR_SETUP);

    auto decode_tile_index = [&](int t, int& nt, int& sp, int& b, int& l0) {
        int mt;
        tile_coords(t, m_tiles, p.band_m, n_tiles * p.split_k, p.split_k, mt, sp, nt);
        b = mt / p.m_tiles_per_b;
        l0 = (mt % p.m_tiles_per_b) * BM;
    };
    auto k_range = [&](int sp, int& kb0, int& kb1) {
        kb0 = (int)((long)p.kb_total * sp / p.split_k);
        kb1 = (int)((long)p.kb_total * (sp + 1) / p.split_k);
    };

    if (warp == 8) {
        // ===================== TMA producer =====================
        // Ring / tap / weight-pointer state advances incrementally (no division or 64-bit multiply per k-block).
        // The whole warp runs the loop on warp-uniform values and one elected lane issues: TMA instructions take their
        // operands from uniform registers, a single divergent thread pays an ELECT + R2UR waterfall per instruction.
        if (SLAB) {
            uint32_t s = 0, ph = 0;
            const size_t sub_stride = (size_t)p.kb_total * W_SUB_BYTES;
            const uint32_t panel_bytes = (uint32_t)p.slab_rows * 128u;
            int it = 0;
            for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++it) {
                int nt, sp, b, l0, kb0, kb1;
                decode_tile_index(t, nt, sp, b, l0);
                k_range(sp, kb0, kb1);
                const int buf = it & 1;
                ptx::mbar_wait(&raw_empty[buf], ((it >> 1) & 1) ^ 1);
                if (ptx::elect_one()) {
                    ptx::mbar_expect_tx(&raw_full[buf], panel_bytes * (uint32_t)p.kb_per_tap);
                    for (int pnl = 0; pnl < p.kb_per_tap; ++pnl)
                        ptx::tma_load_3d(sA + buf * SLAB_BYTES + pnl * SLAB_PANEL_BYTES, &tmA, &raw_full[buf], pnl * BK, l0 + p.tap_offset, b);
                }
                __syncwarp();
                const int subs_valid = min(SUBS, p.n_tiles128 - nt * SUBS);
                const uint8_t* wptr = p.W + (size_t)(nt * SUBS) * sub_stride + (size_t)kb0 * W_SUB_BYTES;
                // weight tiles of TWO k-blocks per ring slot: one full/empty handshake per eight MMAs (the ~440 cycles of issue-side
                // latency of a handshake do not hide under four 64-cycle N=128 MMAs, profiles/r1_probe_mma_handshake.log)
                for (int kb = kb0; kb < kb1; kb += 2) {
                    const int n = min(2, kb1 - kb);
                    ptx::mbar_wait(&empty[s], ph ^ 1);
                    if (ptx::elect_one()) {
                        ptx::mbar_expect_tx(&full[s], (uint32_t)(n * subs_valid * W_SUB_BYTES));
                        for (int i = 0; i < n; ++i)
                            for (int sub = 0; sub < subs_valid; ++sub)
                                ptx::bulk_load(sB + (2 * s + i) * SL::B_STAGE + sub * W_SUB_BYTES, wptr + (size_t)i * W_SUB_BYTES + sub * sub_stride,
                                               W_SUB_BYTES, &full[s]);
                    }
                    __syncwarp();
                    wptr += (size_t)n * W_SUB_BYTES;
                    if (++s == (uint32_t)(STAGES / 2)) {
                        s = 0;
                        ph ^= 1;
                    }
                }
            }
        } else {
            uint32_t s = 0, ph = 0;
            const size_t sub_stride = (size_t)p.kb_total * (QUANT ? (size_t)p.tile_bytes : (size_t)W_SUB_BYTES);
            for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
                int nt, sp, b, l0, kb0, kb1;
                decode_tile_index(t, nt, sp, b, l0);
                k_range(sp, kb0, kb1);
                const int subs_valid = min(SUBS, p.n_tiles128 - nt * SUBS);
                int tap = kb0 / p.kb_per_tap;
                int kc = kb0 - tap * p.kb_per_tap;
                int row = l0 + tap * p.tap_stride + p.tap_offset;
                const uint8_t* wptr = p.W + (size_t)(nt * SUBS) * sub_stride + (size_t)kb0 * (QUANT ? (size_t)p.tile_bytes : (size_t)W_SUB_BYTES);
                for (int kb = kb0; kb < kb1; ++kb) {
                    ptx::mbar_wait(&empty[s], ph ^ 1);
                    if (QUANT) {
                        ptx::mbar_wait(&raw_empty[s], ph ^ 1);
                        if (ptx::elect_one()) {
                            ptx::mbar_expect_tx(&full[s], A_STAGE_BYTES);
                            ptx::tma_load_3d(sA + s * A_STAGE_BYTES, &tmA, &full[s], kc * BK, row, b);
                            ptx::mbar_expect_tx(&raw_full[s], (uint32_t)(subs_valid * p.tile_bytes));
                            for (int sub = 0; sub < subs_valid; ++sub)
                                ptx::bulk_load(sR + s * SL::RAW_STAGE + sub * RAW_MAX_BYTES, wptr + sub * sub_stride, p.tile_bytes, &raw_full[s]);
                        }
                        wptr += p.tile_bytes;
                    } else {
                        if (ptx::elect_one()) {
                            ptx::mbar_expect_tx(&full[s], A_STAGE_BYTES + subs_valid * W_SUB_BYTES);
                            ptx::tma_load_3d(sA + s * A_STAGE_BYTES, &tmA, &full[s], kc * BK, row, b);
                            for (int sub = 0; sub < subs_valid; ++sub)
                                ptx::bulk_load(sB + s * SL::B_STAGE + sub * W_SUB_BYTES, wptr + sub * sub_stride, W_SUB_BYTES, &full[s]);
                        }
                        wptr += W_SUB_BYTES;
                    }
                    __syncwarp();
                    if (++kc == p.kb_per_tap) {
                        kc = 0;
                        row += p.tap_stride;
                    }
                    if (++s == (uint32_t)STAGES) {
                        s = 0;
                        ph ^= 1;
                    }
                }
            }
        }
    } else if (warp == 9) {
        // ===================== MMA issuer =====================
        // The WHOLE warp runs the loop on warp-uniform values (ring state, descriptors, TMEM address — the latter made
        // provably uniform with a shuffle) and one elected lane issues.  tcgen05.mma / commit take their operands from
        // UNIFORM registers: when a single divergent thread ran this loop every MMA was wrapped in an ELECT + 5x R2UR
        // waterfall (~150 cycles per MMA, 650 per k-block against 256 of tensor-pipe time: profiles/r1_vae_conv_ncu.md).
        {
            constexpr uint32_t idesc = ptx::umma_idesc_f16(BM, BN);
            const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
            const uint64_t desc_a0 = ptx::umma_desc_sw128(ptx::smem_u32(sA)), desc_b0 = ptx::umma_desc_sw128(ptx::smem_u32(sB));
            uint32_t s = 0, ph = 0;
            bool first = true;
            int it = 0;
            for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++it) {
                int nt, sp, b, l0, kb0, kb1;
                decode_tile_index(t, nt, sp, b, l0);
                k_range(sp, kb0, kb1);
                const int a = it & 1;
                const uint32_t aph = (it >> 1) & 1;
                if (lane == 0 && it == 8) trace_stamp(p, TR_T8_MMA_WAIT);
                ptx::mbar_wait(&acc_empty[a], aph ^ 1);
                if (SLAB) ptx::mbar_wait(&raw_full[it & 1], (it >> 1) & 1);  // this tile's activation slab has landed
                ptx::tc_fence_after();
                if (lane == 0 && it == 8) trace_stamp(p, TR_T8_MMA_GO);
                const uint32_t d_tmem = tmem_u + (uint32_t)(a * BN);
                uint32_t acc_flag = 0;
                // SLAB: descriptor of (panel kc, row offset tap * dilation) inside the tile's slab, advanced incrementally
                uint32_t slab_off = (uint32_t)((it & 1) * (SLAB_BYTES >> 4)) + (uint32_t)((kb0 / p.kb_per_tap) * p.tap_stride * 8);
                int kc = kb0 % p.kb_per_tap;
                if (SLAB) {  // ring slot = two k-blocks of weight tiles, A from the slab
                    for (int kb = kb0; kb < kb1; kb += 2) {
                        const int n = min(2, kb1 - kb);
                        ptx::mbar_wait(&full[s], ph);
                        ptx::tc_fence_after();
                        if (first) {
                            if (lane == 0) trace_stamp(p, TR_FIRST_FULL);
                            first = false;
                        }
                        uint64_t da2[2];
#pragma unroll
                        for (int i = 0; i < 2; ++i) {
                            da2[i] = desc_a0 + (uint64_t)(slab_off + (uint32_t)kc * (SLAB_PANEL_BYTES >> 4));
                            if (++kc == p.kb_per_tap) {
                                kc = 0;
                                slab_off += (uint32_t)(p.tap_stride * 8);
                            }
                        }
                        const uint64_t db = desc_b0 + (uint64_t)(2 * s * (SL::B_STAGE >> 4));
                        if (ptx::elect_one()) {
#pragma unroll
                            for (int k = 0; k < BK / 16; ++k) ptx::umma_f16(d_tmem, da2[0] + 2 * k, db + 2 * k, idesc, k > 0 ? 1u : acc_flag);
                            if (n > 1) {
#pragma unroll
                                for (int k = 0; k < BK / 16; ++k) ptx::umma_f16(d_tmem, da2[1] + 2 * k, db + (SL::B_STAGE >> 4) + 2 * k, idesc, 1u);
                            }
                            ptx::umma_commit(&empty[s]);
                        }
                        __syncwarp();
                        acc_flag = 1;
                        if (++s == (uint32_t)(STAGES / 2)) {
                            s = 0;
                            ph ^= 1;
                        }
                    }
                } else {
                long long waited = 0;
                for (int kb = kb0; kb < kb1; ++kb) {
                    const bool timed = p.trace != nullptr && it == 8;
                    const long long tw = timed ? clock64() : 0;
                    ptx::mbar_wait(&full[s], ph);
                    ptx::tc_fence_after();
                    if (timed) {
                        waited += clock64() - tw;
                        if (lane == 0) p.trace[(size_t)blockIdx.x * TR_N + 15] = waited;  // cycles waiting for operands in the 9th tile
                    }
                    if (first) {
                        if (lane == 0) trace_stamp(p, TR_FIRST_FULL);
                        first = false;
                    }
                    const uint64_t da = SLAB ? desc_a0 + (uint64_t)(slab_off + (uint32_t)kc * (SLAB_PANEL_BYTES >> 4))
                                             : desc_a0 + (uint64_t)(s * (A_STAGE_BYTES >> 4));
                    const uint64_t db = desc_b0 + (uint64_t)(s * (SL::B_STAGE >> 4));
                    if (SLAB && ++kc == p.kb_per_tap) {
                        kc = 0;
                        slab_off += (uint32_t)(p.tap_stride * 8);  // next tap: `dilation` rows of 128 bytes further down
                    }
                    if (ptx::elect_one()) {
#pragma unroll
                        for (int k = 0; k < BK / 16; ++k) ptx::umma_f16(d_tmem, da + 2 * k, db + 2 * k, idesc, k > 0 ? 1u : acc_flag);
                        ptx::umma_commit(&empty[s]);
                    }
                    __syncwarp();
                    acc_flag = 1;
                    if (++s == (uint32_t)STAGES) {
                        s = 0;
                        ph ^= 1;
                    }
                }
                }
                if (ptx::elect_one()) {
                    ptx::umma_commit(&acc_full[a]);
                    if (SLAB) ptx::umma_commit(&raw_empty[it & 1]);  // every MMA reading this slab has completed
                }
                __syncwarp();
                if (lane == 0) {
                    trace_stamp(p, TR_LAST_MMA);
                    if (it == 8) trace_stamp(p, TR_T8_MMA_DONE);
                }
            }
        }
    } else if (QUANT && warp < 4) {
        // ===================== weight decoders =====================
        {
            const int row = threadIdx.x;  // 0..127
            uint32_t g = 0;
            for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
                int nt, sp, b, l0, kb0, kb1;
                decode_tile_index(t, nt, sp, b, l0);
                k_range(sp, kb0, kb1);
                const int subs_valid = min(SUBS, p.n_tiles128 - nt * SUBS);
                for (int kb = kb0; kb < kb1; ++kb, ++g) {
                    const int s = g % STAGES;
                    const uint32_t ph = (g / STAGES) & 1;
                    ptx::mbar_wait(&raw_full[s], ph);
                    ptx::mbar_wait(&empty[s], ph ^ 1);
                    for (int sub = 0; sub < subs_valid; ++sub)
                        decode_tile(p.wtype, sR + s * SL::RAW_STAGE + sub * RAW_MAX_BYTES,
                                    sB + s * SL::B_STAGE + sub * W_SUB_BYTES, row);
                    ptx::fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        ptx::mbar_arrive(&full[s]);
                        ptx::mbar_arrive(&raw_empty[s]);
                    }
                }
            }
        }
    } else {
        // ===================== epilogue =====================
        // warp w may touch TMEM lanes 32*(w%4)..+31.  QUANT: warps 4..7, one group.  F16 weights: warps 0..3 have no
        // decoding to do and form a second group; the two groups take alternate column chunks / heads of a tile.
        const int q = warp & 3;
        const int part = QUANT ? 0 : (warp >> 2);
        constexpr int NPARTS = QUANT ? 1 : 2;
        int it = 0;
        // Warm-up pass: with one or two tiles per CTA the epilogue is straight-line code that runs once, at instruction-fetch
        // speed (ncu: `no_inst` stalls all over it).  The epilogue warps are idle while the first tile's MMAs run, so they
        // execute the SAME code once as a dry run (loop iteration with `dry` set: garbage accumulator, real loads, stores
        // predicated off) and the real pass finds it in the instruction cache.
        bool dry = WARM;
        for (int t = blockIdx.x; t < total_tiles;) {
            int nt, sp, b, l0, kb0, kb1;
            decode_tile_index(t, nt, sp, b, l0);
            k_range(sp, kb0, kb1);
            const int a = it & 1;
            const uint32_t aph = (it >> 1) & 1;
            constexpr int EPI_THREADS = SL::EPI_WARPS * 32;
            if (!dry) {
                if (p.epi.mode == EPI_STORE) {  // column terms of this tile, fetched under the tile's MMAs
                    epi_bar_sync(EPI_THREADS);  // every warp is done reading the previous tile's terms
                    stage_column_terms<BN, EPI_THREADS>(p, b, nt * BN, sp == 0, QUANT ? (int)threadIdx.x - 128 : (int)threadIdx.x, sCol, sCol + BN, sCol + 2 * BN);
                    epi_bar_sync(EPI_THREADS);
                } else if (p.epi.mode == EPI_HEADNORM) {
                    epi_bar_sync(EPI_THREADS);
                    stage_head_weights<BN, EPI_THREADS>(p, nt * BN, QUANT ? (int)threadIdx.x - 128 : (int)threadIdx.x, sCol);
                    epi_bar_sync(EPI_THREADS);
                }
                if (warp == 4 && lane == 0 && it == 8) trace_stamp(p, TR_T8_EPI_WAIT);
                ptx::mbar_wait(&acc_full[a], aph);
                ptx::tc_fence_after();
                if (warp == 4 && lane == 0) trace_stamp(p, TR_ACC_READY);
                if (warp == 4 && lane == 0 && it == 8) trace_stamp(p, TR_T8_EPI_GO);
            }
            const int l = l0 + q * 32 + lane;
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(a * BN);
            if (p.epi.mode == EPI_HEADNORM) {
                float* stg = sStg + (QUANT ? q : warp) * (STG_WARP_BYTES / 4);
                if (NPARTS == 2 && BN == 128) {  // a single head per tile: the two warp groups split it by chunk pair
                    const int n0 = nt * BN;
                    if (n0 < p.N && kb1 > kb0) epilogue_head128<WARM>(p, taddr, b, !dry, l0 + q * 32, lane, n0, stg, sCol, part, 2);
                } else {
#pragma unroll 1
                    for (int hh = part; hh < BN / 128; hh += NPARTS) {
                        const int n0 = nt * BN + hh * 128;
                        if (n0 >= p.N) break;  // warp-uniform
                        if (kb1 > kb0) epilogue_head128<WARM>(p, taddr + hh * 128, b, !dry, l0 + q * 32, lane, n0, stg, sCol + hh * 128, 0, 1);
                    }
                }
            } else
#pragma unroll 1
            for (int c = part; c < BN / 32; c += NPARTS) {
                const int n0 = nt * BN + c * 32;
                if (n0 >= p.N) break;  // warp-uniform
                uint32_t r[32];
                ptx::tmem_ld32(taddr + c * 32, r);
                ptx::tmem_ld_wait();
                if (warp == 4 && lane == 0 && c == part) trace_stamp(p, TR_EPI_LD0);
                float acc[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) acc[i] = __uint_as_float(r[i]);
                if (kb1 > kb0) {
                    float* stg = sStg + (QUANT ? q : warp) * (STG_WARP_BYTES / 4);
                    if (p.epi.mode != EPI_STORE || !epilogue_store_staged<WARM>(p, &tmOut, acc, b, l0 + q * 32, lane, n0, sCol + c * 32, sCol + BN + c * 32, stg, dry,
                                                                                sCol + 2 * BN + c * 32, sCol + 3 * BN + c * 32))
                        epilogue_row32<WARM>(p, acc, b, l, n0, sp == 0, dry);
                }
                if (warp == 4 && lane == 0 && c == part) trace_stamp(p, TR_EPI_CHUNK0);
            }
            if (dry) {
                dry = false;
                continue;  // same tile again, for real
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&acc_empty[a]);
            if (warp == 4 && lane == 0) trace_stamp(p, TR_EPI_DONE);
            if (warp == 4 && lane == 0 && it == 8) trace_stamp(p, TR_T8_EPI_DONE);
            t += gridDim.x;
            ++it;
        }
    }

    if (p.tma_reduce) ptx::bulk_wait0();  // every reduce-add this thread issued has been performed
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 9) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<TMEM_COLS>(tmem_base);
    }
    if (threadIdx.x == 0) trace_stamp(p, TR_END);
}

// ---------------------------------------------------------------- CTA-pair kernel (cta_group::2), F16 weight tiles
// A cluster of two CTAs (one TPC) computes a 256 x 256 tile with M = 256 MMAs: CTA r owns rows [l0 + 128 r, +128) of A and
// of the accumulator (its own TMEM) and stages only HALF of the weight tile (rows n0 + 128 r ..) — the tensor core reads
// the other half from the peer's shared memory.  Per k-block a CTA pulls 16 KB (A) + 16 KB (W) from L2 instead of 16 + 32,
// which is what bounds the single-CTA 128 x 256 tile (L2 -> SM operand delivery, profiles/r1_gemm_ncu_big1.md).
// Synchronisation: every CTA has local full/empty/acc_full barriers; CTA 1 forwards "my stage landed" to the leader with
// a remote mbarrier arrive (peer_full), the leader's tcgen05.commit multicasts "stage consumed" / "accumulator ready" to
// both CTAs, and both CTAs' epilogue warps arrive on the leader's acc_empty.
constexpr int STAGES2 = 6;
constexpr int STAGE2_BYTES = A_STAGE_BYTES + W_SUB_BYTES;  // 32 KB
constexpr int SMEM2_BYTES = STAGES2 * STAGE2_BYTES + 8 * STG_WARP_BYTES + 2 * 256 * 4 + 512;
// QUANT variant (packed Q8_0 / Q4_K / Q6_K tiles decoded in shared memory): the pair shape is what makes in-kernel decoding
// affordable at large M — a CTA decodes ONE 128 x 64 sub-tile per k-block (its half of the weight tile) while the pair's
// tensor cores spend 512 cycles on it (M = 256 x N = 256), instead of two sub-tiles per 512 cycles in the single-CTA kernel,
// whose 3-stage ring (48 KB + 17 KB of codes per stage) left the MMA thread waiting 966 of 1444 cycles per k-block
// (profiles/r2_quant_trace_d.log).
// Warp budget of the QUANT variants: 14 warps = 12 workers + TMA + MMA.  DEC_GROUPS = 2 (default): 4 epilogue warps + two decoder
// groups on alternate k-blocks — one group of 4 warps does not finish a 128 x 64 sub-tile inside the 512-cycle k-block (measured);
// DEC_GROUPS = 1: 8 epilogue warps + one decoder group (A/B variant).  NSQ = ring depth: the packed-code slots are sized by the weight type at run time
// (Q4_K 5120 B, Q6_K 7168 B, Q8_0 8704 B), so Q4_K affords 5 stages next to 8 transpose buffers, the others 4.
inline __host__ __device__ int raw2_stride(int tile_bytes) { return (tile_bytes + 127) & ~127; }
inline __host__ __device__ int raw2_total(int ns, int tile_bytes) { return (ns * raw2_stride(tile_bytes) + 1023) / 1024 * 1024; }
// pub: decoders publish their rows through one 16 KB staging buffer per decoder group and the bulk-copy engine (see gemm_tc2_kernel)
inline int smem2q_bytes(int ns, int epi_warps, int tile_bytes, bool pub = false) {
    return ns * STAGE2_BYTES + raw2_total(ns, tile_bytes) + epi_warps * STG_WARP_BYTES + 2 * 256 * 4 + 512 + (pub ? 2 * W_SUB_BYTES : 0);
}
// EPIW = 8 with two decoder groups makes 18 warps (576 threads, 112 registers per thread: the epilogue spills a little more, but
// twice as many warps share it — the K = 2048 shapes with heavy epilogues (QKV head-norm, SwiGLU) were epilogue-bound with 4)
constexpr int tc2q_threads(int epiw, int dg) { return (epiw + 4 * dg + 2) * 32; }

template <bool WARM, bool QUANT, int NSQ = 5, int DEC_GROUPS = 2, int EPIW = 4>
__global__ void __launch_bounds__(QUANT ? tc2q_threads(EPIW, DEC_GROUPS) : NUM_THREADS, 1)
gemm_tc2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmOut, const KParams p) {
    constexpr int BN = 256;
    constexpr int TMEM_COLS = 2 * BN;
    constexpr int NS = QUANT ? NSQ : STAGES2;
    constexpr int EPI_WARPS = QUANT ? EPIW : 8;
    constexpr int NPARTS = EPI_WARPS / 4;
    constexpr int DEC0 = EPI_WARPS;  // first decoder warp
    constexpr int TMA_WARP = QUANT ? EPIW + 4 * DEC_GROUPS : 8;
    constexpr int MMA_WARP = TMA_WARP + 1;
    const int raw_stride = QUANT ? raw2_stride(p.tile_bytes) : 0;
    extern __shared__ __align__(1024) uint8_t smem[];
    if (threadIdx.x == 0 && (ptx::smem_u32(smem) & 1023u) != 0) __trap();
    uint8_t* sA = smem;                       // [NS][16 KB]
    uint8_t* sB = sA + NS * A_STAGE_BYTES;    // [NS][16 KB]  this CTA's half of the weight tile (fp16 UMMA image)
    uint8_t* sR = sB + NS * W_SUB_BYTES;      // QUANT: [NS][raw_stride] packed codes as they land
    constexpr int nraw = NS;   // the packed-code ring is as deep as the stage ring
    float* sStg = reinterpret_cast<float*>(sR + (QUANT ? raw2_total(nraw, p.tile_bytes) : 0));  // [EPI_WARPS][32 x 32 fp32] epilogue transpose buffers (1 KB aligned:
                                                                          // they are the 128-byte-swizzle source image of the TMA reduce-add boxes)
    float* sCol = sStg + EPI_WARPS * (STG_WARP_BYTES / 4);  // [2][256]: bias, gate of the current tile
    uint64_t* bars = reinterpret_cast<uint64_t*>(sCol + 2 * 256);
    uint64_t* full = bars;               // [S] local operands landed (QUANT: A landed + this CTA's sub-tile decoded)
    uint64_t* empty = bars + NS;         // [S] MMAs reading this CTA's stage completed (multicast commit)
    uint64_t* peer_full = bars + 2 * NS; // [S] leader only: CTA 1's stage landed
    uint64_t* raw_full = bars + 3 * NS;  // [S] QUANT: packed codes landed
    uint64_t* raw_empty = bars + 4 * NS; // [S] QUANT: decoders done with the packed codes
    uint64_t* acc_full = bars + 5 * NS;  // [2] local (multicast commit)
    uint64_t* acc_empty = acc_full + 2;  // [2] leader only: the epilogue warps of the pair
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
    uint8_t* sDec = reinterpret_cast<uint8_t*>(bars) + 512;   // QUANT + qpub: [DEC_GROUPS][16 KB] decoded rows on their way to sB
    const bool pub = QUANT && p.qpub != 0;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = ptx::cluster_ctarank();
    const int cluster_id = blockIdx.x >> 1;
    const int n_clusters = gridDim.x >> 1;

    const int m_pairs_per_b = (p.l_out + 2 * BM - 1) / (2 * BM);
    const int m_tiles = p.B * m_pairs_per_b;
    const int n_tiles = p.n_tiles128 >> 1;  // launcher guarantees an even count
    const int total_tiles = n_tiles * p.split_k * m_tiles;

    if (threadIdx.x == 0) {
        for (int s = 0; s < NS; ++s) {
            ptx::mbar_init(&full[s], (QUANT && !pub) ? 1 + 4 : 1);
            ptx::mbar_init(&empty[s], 1);
            ptx::mbar_init(&peer_full[s], 1);
            ptx::mbar_init(&raw_full[s], 1);
            ptx::mbar_init(&raw_empty[s], 4);
        }
        for (int a = 0; a < 2; ++a) {
            ptx::mbar_init(&acc_full[a], 1);
            ptx::mbar_init(&acc_empty[a], 2 * EPI_WARPS);
        }
        ptx::fence_mbar_init();
    }
    if (warp == MMA_WARP) ptx::tmem_alloc_2cta<TMEM_COLS>(tmem_slot);
    if (warp == TMA_WARP && lane == 0) ptx::prefetch_tmap(&tmA);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    ptx::pdl_wait();
    ptx::pdl_launch();

    auto decode_tile_index = [&](int t, int& nt, int& sp, int& b, int& l0) {
        int mt;
        tile_coords(t, m_tiles, p.band_m, n_tiles * p.split_k, p.split_k, mt, sp, nt);
        b = mt / m_pairs_per_b;
        l0 = (mt % m_pairs_per_b) * (2 * BM) + (int)rank * BM;  // this CTA's 128 rows
    };
    auto k_range = [&](int sp, int& kb0, int& kb1) {
        kb0 = (int)((long)p.kb_total * sp / p.split_k);
        kb1 = (int)((long)p.kb_total * (sp + 1) / p.split_k);
    };

    if (warp == TMA_WARP) {
        // ===================== TMA producer (both CTAs): whole warp, elected lane issues =====================
        {
            uint32_t s = 0, ph = 0;
            uint32_t rs = 0, rph = 0;   // packed-code ring (QUANT)
            const size_t w_tile = QUANT ? (size_t)p.tile_bytes : (size_t)W_SUB_BYTES;
            const size_t sub_stride = (size_t)p.kb_total * w_tile;
            for (int t = cluster_id; t < total_tiles; t += n_clusters) {
                int nt, sp, b, l0, kb0, kb1;
                decode_tile_index(t, nt, sp, b, l0);
                k_range(sp, kb0, kb1);
                int tap = kb0 / p.kb_per_tap;
                int kc = kb0 - tap * p.kb_per_tap;
                int row = l0 + tap * p.tap_stride + p.tap_offset;
                const uint8_t* wptr = p.W + (size_t)(nt * 2 + (int)rank) * sub_stride + (size_t)kb0 * w_tile;
                for (int kb = kb0; kb < kb1; ++kb) {
                    ptx::mbar_wait(&empty[s], ph ^ 1);
                    if (QUANT) {
                        ptx::mbar_wait(&raw_empty[rs], rph ^ 1);
                        if (ptx::elect_one()) {
                            ptx::mbar_expect_tx(&full[s], pub ? A_STAGE_BYTES + W_SUB_BYTES : A_STAGE_BYTES);
                            ptx::tma_load_3d(sA + s * A_STAGE_BYTES, &tmA, &full[s], kc * BK, row, b);
                            ptx::mbar_expect_tx(&raw_full[rs], (uint32_t)p.tile_bytes);
                            ptx::bulk_load(sR + rs * raw_stride, wptr, p.tile_bytes, &raw_full[rs]);
                        }
                        if (++rs == (uint32_t)nraw) {
                            rs = 0;
                            rph ^= 1;
                        }
                    } else {
                        if (ptx::elect_one()) {
                            ptx::mbar_expect_tx(&full[s], STAGE2_BYTES);
                            ptx::tma_load_3d(sA + s * A_STAGE_BYTES, &tmA, &full[s], kc * BK, row, b);
                            ptx::bulk_load(sB + s * W_SUB_BYTES, wptr, W_SUB_BYTES, &full[s]);
                        }
                    }
                    __syncwarp();
                    wptr += w_tile;
                    if (++kc == p.kb_per_tap) {
                        kc = 0;
                        row += p.tap_stride;
                    }
                    if (++s == (uint32_t)NS) {
                        s = 0;
                        ph ^= 1;
                    }
                }
            }
        }
    } else if (warp == MMA_WARP) {
        if (rank == 0) {
            // ===================== MMA issuer (leader CTA): whole warp on uniform values, elected lane issues =====================
            constexpr uint32_t idesc = ptx::umma_idesc_f16(2 * BM, BN);
            const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
            const uint64_t desc_a0 = ptx::umma_desc_sw128(ptx::smem_u32(sA)), desc_b0 = ptx::umma_desc_sw128(ptx::smem_u32(sB));
            uint32_t s = 0, ph = 0;
            int it = 0;
            for (int t = cluster_id; t < total_tiles; t += n_clusters, ++it) {
                int nt, sp, b, l0, kb0, kb1;
                decode_tile_index(t, nt, sp, b, l0);
                k_range(sp, kb0, kb1);
                const int a = it & 1;
                const uint32_t aph = (it >> 1) & 1;
                ptx::mbar_wait(&acc_empty[a], aph ^ 1);
                ptx::tc_fence_after();
                const uint32_t d_tmem = tmem_u + (uint32_t)(a * BN);
                uint32_t acc_flag = 0;
                for (int kb = kb0; kb < kb1; ++kb) {
                    ptx::mbar_wait(&full[s], ph);
                    ptx::mbar_wait(&peer_full[s], ph);
                    ptx::tc_fence_after();
                    const uint64_t da = desc_a0 + (uint64_t)(s * (A_STAGE_BYTES >> 4)), db = desc_b0 + (uint64_t)(s * (W_SUB_BYTES >> 4));
                    if (ptx::elect_one()) {
#pragma unroll
                        for (int k = 0; k < BK / 16; ++k) ptx::umma_f16_2cta(d_tmem, da + 2 * k, db + 2 * k, idesc, k > 0 ? 1u : acc_flag);
                        ptx::umma_commit_2cta(&empty[s], 3);
                    }
                    __syncwarp();
                    acc_flag = 1;
                    if (++s == (uint32_t)NS) {
                        s = 0;
                        ph ^= 1;
                    }
                }
                if (ptx::elect_one()) ptx::umma_commit_2cta(&acc_full[a], 3);
                __syncwarp();
            }
        } else {
            // ===================== relay (CTA 1): local "stage landed" -> leader's peer_full =====================
            if (lane == 0) {
                uint32_t s = 0, ph = 0;
                for (int t = cluster_id; t < total_tiles; t += n_clusters) {
                    int nt, sp, b, l0, kb0, kb1;
                    decode_tile_index(t, nt, sp, b, l0);
                    k_range(sp, kb0, kb1);
                    for (int kb = kb0; kb < kb1; ++kb) {
                        ptx::mbar_wait(&full[s], ph);
                        ptx::mbar_arrive_cluster(ptx::mapa(&peer_full[s], 0));
                        if (++s == (uint32_t)NS) {
                            s = 0;
                            ph ^= 1;
                        }
                    }
                }
            }
        }
    } else if (QUANT && warp >= DEC0) {
        // ===================== weight decoders (both CTAs): DEC_GROUPS groups of 4 warps on alternate k-blocks of the ring =====
        // thread = one row of the CTA's 128 x 64 sub-tile; the decoded tile is written with generic stores, published to the
        // tensor-core (async) proxy with fence.proxy.async, then counted on full[s] next to the TMA's A box.
        const int grp = (warp - DEC0) >> 2;
        const int row = (int)threadIdx.x - DEC0 * 32 - grp * 128;  // 0..127
        uint32_t g = 0;
        // PUBLISH mode (p.qpub; even ring depth): the rows are decoded into a staging buffer of the group and each warp hands its 4 KB
        // (32 rows of the 128-byte-swizzle image, contiguous) to the bulk-copy engine, which writes them into sB[s] and counts the bytes
        // on full[s] (complete_tx) next to the TMA's A box.  sB is then written by the async proxy only, exactly like the F16 kernel's
        // weight tiles, so "full[s] completed" means "every byte is in shared memory" for BOTH tensor cores of the pair, and the
        // writers' fence only has to cover the CTA's own copy engine (fence.proxy.async.shared::cta).  The staging buffer is free again
        // when the copy out of it has completed = the previous stage of this group is full.
        uint8_t* myDec = sDec + grp * W_SUB_BYTES;
        const int wq = (warp - DEC0) & 3;
        bool have_prev = false;
        uint32_t prev_s = 0, prev_ph = 0;
        for (int t = cluster_id; t < total_tiles; t += n_clusters) {
            int nt, sp, b, l0, kb0, kb1;
            decode_tile_index(t, nt, sp, b, l0);
            k_range(sp, kb0, kb1);
            for (int kb = kb0; kb < kb1; ++kb, ++g) {
                if (DEC_GROUPS == 2 && (int)(g & 1u) != grp) continue;
                const uint32_t s = g % (uint32_t)NS;
                const uint32_t ph = (g / (uint32_t)NS) & 1u;
                const uint32_t rs = g % (uint32_t)nraw;
                const uint32_t rph = (g / (uint32_t)nraw) & 1u;
                ptx::mbar_wait(&raw_full[rs], rph);
                if (pub) {
                    uint4 out[8];   // the row is decoded into registers while the previous copy still reads the staging buffer
                    decode_row(p.wtype, sR + rs * raw_stride, row, out);
                    uint8_t* dec = myDec;
                    if (have_prev) ptx::mbar_wait(&full[prev_s], prev_ph);   // the previous copy has read the staging buffer
                    store_row(dec, row, out);
                    ptx::fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        ptx::mbar_arrive(&raw_empty[rs]);
                        ptx::mbar_wait(&empty[s], ph ^ 1);   // only the copy INTO sB[s] has to wait for the MMAs that read its old contents
                        ptx::bulk_copy_s2s(sB + s * W_SUB_BYTES + wq * 4096, dec + wq * 4096, 4096, &full[s]);
                    }
                    __syncwarp();
                    have_prev = true;
                    prev_s = s;
                    prev_ph = ph;
                    continue;
                }
                ptx::mbar_wait(&empty[s], ph ^ 1);
                decode_tile(p.wtype, sR + rs * raw_stride, sB + s * W_SUB_BYTES, row);
                // The decoded sub-tile is read by BOTH tensor cores of the pair (each CTA stages half of the 256-row weight tile): the peer's
                // reads come through the shared::cluster window, so the writers' proxy fence must cover that window.  With the CTA-scoped
                // fence (fence.proxy.async.shared::cta = MEMBAR.ALL.CTA + FENCE.VIEW.ASYNC.S) a tensor core occasionally consumed a
                // sub-tile that was not completely visible yet: one 256-row tile off by ~1e-3, 8 distinct results in 12 identical
                // calls at M = 48 000 (scripts/probes/determinism.py, profiles/r2_determinism_*.log).  The cluster-scoped fence compiles to
                // MEMBAR.ALL.GPU + FENCE.VIEW.ASYNC.S and costs ~25 % of the kernel's throughput (profiles/r2_quant_sweep_fence_s8.log);
                // no cheaper placement was found (one lane per warp, the relay, the leader: same cost).
                // ACE_B200_GEMM_DBG bit 1 restores the CTA-scoped fence for A/B measurements.
                if (p.dbg & 1) ptx::fence_proxy_async_smem();
                else asm volatile("fence.proxy.async.shared::cluster;" ::: "memory");
                __syncwarp();
                if (lane == 0) {
                    ptx::mbar_arrive(&full[s]);
                    ptx::mbar_arrive(&raw_empty[rs]);
                }
            }
        }
    } else {
        // ===================== epilogue: warps 0..EPI_WARPS-1 of both CTAs, NPARTS groups on alternate column chunks ==========
        const int q = warp & 3;
        const int part = warp >> 2;
        constexpr int EPI_THREADS = EPI_WARPS * 32;
        int it = 0;
        bool dry = WARM;  // instruction-cache warm-up pass, see gemm_tc_kernel
        for (int t = cluster_id; t < total_tiles;) {
            int nt, sp, b, l0, kb0, kb1;
            decode_tile_index(t, nt, sp, b, l0);
            k_range(sp, kb0, kb1);
            const int a = it & 1;
            const uint32_t aph = (it >> 1) & 1;
            if (!dry) {
                if (p.epi.mode == EPI_STORE) {
                    epi_bar_sync(EPI_THREADS);
                    stage_column_terms<BN, EPI_THREADS>(p, b, nt * BN, sp == 0, (int)threadIdx.x, sCol, sCol + BN);
                    epi_bar_sync(EPI_THREADS);
                } else if (p.epi.mode == EPI_HEADNORM) {
                    epi_bar_sync(EPI_THREADS);
                    stage_head_weights<BN, EPI_THREADS>(p, nt * BN, (int)threadIdx.x, sCol);
                    epi_bar_sync(EPI_THREADS);
                }
                ptx::mbar_wait(&acc_full[a], aph);
                ptx::tc_fence_after();
            }
            const int l = l0 + q * 32 + lane;
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(a * BN);
            if (p.epi.mode == EPI_HEADNORM) {
#pragma unroll 1
                for (int hh = part; hh < BN / 128; hh += NPARTS) {
                    const int n0 = nt * BN + hh * 128;
                    if (n0 >= p.N) break;
                    if (kb1 > kb0) epilogue_head128<WARM>(p, taddr + hh * 128, b, !dry, l0 + q * 32, lane, n0, sStg + warp * (STG_WARP_BYTES / 4), sCol + hh * 128, 0, 1);
                }
            } else
#pragma unroll 1
            for (int c = part; c < BN / 32; c += NPARTS) {
                const int n0 = nt * BN + c * 32;
                if (n0 >= p.N) break;
                uint32_t r[32];
                ptx::tmem_ld32(taddr + c * 32, r);
                ptx::tmem_ld_wait();
                float acc[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) acc[i] = __uint_as_float(r[i]);
                if (kb1 > kb0) {
                    float* stg = sStg + warp * (STG_WARP_BYTES / 4);
                    if (p.epi.mode != EPI_STORE || !epilogue_store_staged<WARM>(p, &tmOut, acc, b, l0 + q * 32, lane, n0, sCol + c * 32, sCol + BN + c * 32, stg, dry))
                        epilogue_row32<WARM>(p, acc, b, l, n0, sp == 0, dry);
                }
            }
            if (dry) {
                dry = false;
                continue;  // same tile again, for real
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive_cluster(ptx::mapa(&acc_empty[a], 0));
            t += n_clusters;
            ++it;
        }
    }

    if (p.tma_reduce) ptx::bulk_wait0();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::cluster_sync();  // the peer's smem / barriers stay alive until every MMA and remote arrive has landed
    if (warp == MMA_WARP) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc_2cta<TMEM_COLS>(tmem_base);
    }
}

// ---------------------------------------------------------------- plain-FMA debug kernel (same operands, same epilogue)
// One thread per (row, 32-column group); decodes the packed weight element-wise. Slow by construction; it exists so
// the tcgen05 kernel can be bisected against an independent implementation on the GPU.
__device__ float packed_weight_at(const KParams& p, int n, int kbt, int k) {
    const int nt = n >> 7, r = n & 127;
    const uint8_t* tile = p.W + ((size_t)nt * p.kb_total + kbt) * p.tile_bytes;
    if (p.wtype == W_F16) {
        const __half h = *reinterpret_cast<const __half*>(tile + r * 128 + (((k >> 3) ^ (r & 7)) << 4) + (k & 7) * 2);
        return __half2float(h);
    }
    if (p.wtype == W_Q8_0) {
        const int8_t c = (int8_t)tile[(k >> 4) * 2048 + r * 16 + (k & 15)];
        const __half s = *reinterpret_cast<const __half*>(tile + 8192 + r * 4 + (k >> 5) * 2);
        return __half2float(__hmul(__int2half_rn(c), s));
    }
    if (p.wtype == W_Q4_K) {
        const int j = k >> 5, wi = (k >> 3) & 3, i = k & 7;
        const uint32_t w = *reinterpret_cast<const uint32_t*>(tile + j * 2048 + r * 16 + wi * 4);
        const int nib = (i >> 1) + ((i & 1) ? 4 : 0);
        const int c = (w >> (4 * nib)) & 0xF;
        const __half ds = *reinterpret_cast<const __half*>(tile + 4096 + r * 8 + j * 4);
        const __half dm = *reinterpret_cast<const __half*>(tile + 4096 + r * 8 + j * 4 + 2);
        return __half2float(__hfma(__int2half_rn(c), ds, __hneg(dm)));
    }
    const int g = k >> 3, i = k & 7;
    const uint32_t w = *reinterpret_cast<const uint32_t*>(tile + (g >> 2) * 2048 + r * 16 + (g & 3) * 4);
    const uint32_t P = *reinterpret_cast<const uint32_t*>(tile + 4096 + r * 16 + (g >> 1) * 4);
    const int nib = (i >> 1) + ((i & 1) ? 4 : 0);
    const int lo = (w >> (4 * nib)) & 0xF;
    const int bit = ((g & 1) ? 8 : 0) + ((i & 1) ? 16 : 0) + 2 * (i >> 1);
    const int hi = (P >> bit) & 3;
    const __half s = *reinterpret_cast<const __half*>(tile + 6144 + r * 8 + (g >> 1) * 2);
    return __half2float(__hmul(__int2half_rn((lo | (hi << 4)) - 32), s));
}

__global__ void gemm_simt_kernel(const __half* A, long a_lstride, long a_bstride, int L, int C, const KParams p) {
    const int groups = (p.N + 31) / 32;
    const long idx = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const long total = (long)p.B * p.l_out * groups;
    if (idx >= total) return;
    const int gq = (int)(idx % groups);
    const long rest = idx / groups;
    const int l = (int)(rest % p.l_out);
    const int b = (int)(rest / p.l_out);
    const int n0 = gq * 32;
    float acc[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) acc[i] = 0.f;
    for (int kbt = 0; kbt < p.kb_total; ++kbt) {
        const int tap = kbt / p.kb_per_tap;
        const int c0 = (kbt - tap * p.kb_per_tap) * BK;
        const int row = l + tap * p.tap_stride + p.tap_offset;
        if (row < 0 || row >= L) continue;
        const __half* a = A + (long)b * a_bstride + (long)row * a_lstride;
        for (int k = 0; k < BK && c0 + k < C; ++k) {
            const float av = __half2float(a[c0 + k]);
            for (int i = 0; i < 32; ++i)
                if (n0 + i < p.N) acc[i] = fmaf(av, packed_weight_at(p, n0 + i, kbt, k), acc[i]);
        }
    }
    KParams q = p;
    q.epi.atomic = 0;  // debug kernel never splits K: plain read-modify-write
    if (p.epi.atomic) q.epi.resid = p.epi.out_f32;
    epilogue_row32<false>(q, acc, b, l, n0, true, false);
}

// ---------------------------------------------------------------- one-off decode of a packed tile to its F16 image
__global__ void __launch_bounds__(128) predecode_kernel(const uint8_t* __restrict__ W, int wtype, int tile_bytes, uint8_t* __restrict__ dst,
                                                        long n_tiles) {
    __shared__ __align__(16) uint8_t raw[RAW_MAX_BYTES];
    __shared__ __align__(16) uint8_t out[W_SUB_BYTES];
    for (long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        const uint4* src = reinterpret_cast<const uint4*>(W + (size_t)t * tile_bytes);
        for (int i = threadIdx.x; i < tile_bytes / 16; i += 128) reinterpret_cast<uint4*>(raw)[i] = src[i];
        __syncthreads();
        decode_tile(wtype, raw, out, threadIdx.x);
        __syncthreads();
        uint4* d = reinterpret_cast<uint4*>(dst + (size_t)t * W_SUB_BYTES);
        for (int i = threadIdx.x; i < W_SUB_BYTES / 16; i += 128) d[i] = reinterpret_cast<const uint4*>(out)[i];
        __syncthreads();
    }
}

}  // namespace

size_t gemm_predecode_bytes(const PackedWeight& w) {
    return (size_t)w.n_tiles * w.taps * w.kb_per_tap * W_SUB_BYTES;
}

cudaError_t gemm_predecode(const PackedWeight& w, void* dst, cudaStream_t stream) {
    if (w.wtype == W_F16 || !w.dev || !dst) return cudaErrorInvalidValue;
    const long n_tiles = (long)w.n_tiles * w.taps * w.kb_per_tap;
    note_launch();
    const int grid = (int)std::min<long>(n_tiles, 148L * 16);
    predecode_kernel<<<grid, 128, 0, stream>>>(static_cast<const uint8_t*>(w.dev), w.wtype, w.tile_bytes, static_cast<uint8_t*>(dst), n_tiles);
    return cudaGetLastError();
}

// Decode pass: a packed matrix is decoded ONCE per launch into a scratch buffer of the calling stream (<= 50 MB: the largest matrix of
// the DiT), stream-ordered in front of the GEMM, which then runs on F16 tiles.  The weights stay packed in HBM (0.94 GB for the Q4_K
// turbo DiT, no resident twin); the decoded image is written and immediately re-read through the 126 MB L2.  This is the route for
// launches of many rows since the in-kernel decoders of the CTA-pair kernel need a cluster-scoped proxy fence per k-block (see
// gemm_tc2_kernel): gate_up at M = 48 000 runs 941 TFLOP/s with in-kernel decode, the decode pass + F16 kernel costs one ~20 us launch in
// front of a 1.7 ms GEMM.  ACE_B200_DECODE_PASS=0 forbids it (in-kernel decode everywhere), ACE_B200_DECODE_PASS_MIN_M=n moves the row
// threshold (default 2048 rows; below, the GEMMs are weight-stream / latency bound and the in-kernel decode reads 3.6x fewer bytes).
int decode_pass_min_rows() {
    if (const char* e = std::getenv("ACE_B200_DECODE_PASS")) if (e[0] == '0') return 0;   // read per call: tests flip it
    if (const char* e = std::getenv("ACE_B200_DECODE_PASS_MIN_M")) if (e[0]) return std::max(1, std::atoi(e));
    return 2048;
}
namespace {
struct DecodeScratch { void* ptr = nullptr; size_t bytes = 0; };
std::mutex g_scratch_mu;
std::unordered_map<cudaStream_t, DecodeScratch> g_scratch;   // one per stream = per context (contexts never share a stream)
}  // namespace
// called by the owner of `stream` before it destroys the stream (a later stream may reuse the handle value, possibly on another device)
void decode_scratch_release(cudaStream_t stream) {
    std::lock_guard<std::mutex> lk(g_scratch_mu);
    auto it = g_scratch.find(stream);
    if (it == g_scratch.end()) return;
    if (it->second.ptr) {
        cudaStreamSynchronize(stream);
        cudaFree(it->second.ptr);
    }
    g_scratch.erase(it);
}
// scratch of at least `bytes` for launches on `stream` (grown with a synchronising free; released with the stream's context)
void* decode_scratch(cudaStream_t stream, size_t bytes) {
    std::lock_guard<std::mutex> lk(g_scratch_mu);
    DecodeScratch& sc = g_scratch[stream];
    if (sc.bytes < bytes) {
        if (sc.ptr) {
            cudaStreamSynchronize(stream);
            cudaFree(sc.ptr);
        }
        sc.ptr = nullptr;
        sc.bytes = 0;
        if (cudaMalloc(&sc.ptr, bytes) != cudaSuccess) {
            cudaGetLastError();
            sc.ptr = nullptr;
            return nullptr;
        }
        sc.bytes = bytes;
    }
    return sc.ptr;
}

bool predecode_enabled() {  // opt-in since the CTA-pair kernel decodes in shared memory at full speed (read per call: tests flip it)
    const char* e = std::getenv("ACE_B200_PREDECODE");
    return e && e[0] && e[0] != '0';
}

int predecode_min_rows() {
    const char* e = std::getenv("ACE_B200_PREDECODE_MIN_M");  // read per call: tests flip it between launches
    if (e && e[0]) return std::max(1, std::atoi(e));
    return 256;
}

// ---------------------------------------------------------------- host side
namespace {

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

PFN_encodeTiled get_encode_fn() {
    static PFN_encodeTiled fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* sym = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_encodeTiled>(sym);
    });
    return fn;
}

bool make_tmap_a(const GemmProblem& g, CUtensorMap* tm, int box_rows = BM) {
    PFN_encodeTiled enc = get_encode_fn();
    if (!enc) return false;
    const long bstride = g.B > 1 ? g.a_bstride : (long)g.L * g.a_lstride;
    cuuint64_t dims[3] = {(cuuint64_t)g.C, (cuuint64_t)g.L, (cuuint64_t)g.B};
    cuuint64_t strides[2] = {(cuuint64_t)g.a_lstride * 2, (cuuint64_t)bstride * 2};
    cuuint32_t box[3] = {(cuuint32_t)BK, (cuuint32_t)box_rows, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<__half*>(g.A), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS;
}

template <int NSQ, int DG, int EW>
cudaError_t set_tc2q_attr_one() {
    cudaError_t e = cudaFuncSetAttribute(gemm_tc2_kernel<false, true, NSQ, DG, EW>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(gemm_tc2_kernel<true, true, NSQ, DG, EW>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    return e;
}
cudaError_t set_tc2q_attrs() {
    cudaError_t e = set_tc2q_attr_one<5, 1, 8>();
    if (e == cudaSuccess) e = set_tc2q_attr_one<4, 1, 8>();
    if (e == cudaSuccess) e = set_tc2q_attr_one<5, 2, 4>();
    if (e == cudaSuccess) e = set_tc2q_attr_one<4, 2, 4>();
    if (e == cudaSuccess) e = set_tc2q_attr_one<5, 2, 8>();
    if (e == cudaSuccess) e = set_tc2q_attr_one<4, 2, 8>();
    return e;
}
template <int NSQ, int DG, int EW>
cudaError_t launch_tc2q(cudaLaunchConfig_t& cfg, const CUtensorMap& tm, const CUtensorMap& tmo, const KParams& kp) {
    cfg.blockDim = dim3(tc2q_threads(EW, DG));
    return kp.warm_epi ? cudaLaunchKernelEx(&cfg, gemm_tc2_kernel<true, true, NSQ, DG, EW>, tm, tmo, kp)
                       : cudaLaunchKernelEx(&cfg, gemm_tc2_kernel<false, true, NSQ, DG, EW>, tm, tmo, kp);
}

// co-resident CTA pairs on this device (a GPC with an odd SM count cannot pair its last SM); 0 = pairs unavailable
int max_active_pairs() {
    static DeviceOnce once;
    int result = 0;
    once.run([](int& n) {
        n = 0;
        if (cudaFuncSetAttribute(gemm_tc2_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM2_BYTES) == cudaSuccess &&
            cudaFuncSetAttribute(gemm_tc2_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM2_BYTES) == cudaSuccess &&
            set_tc2q_attrs() == cudaSuccess) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(2 * 148);
            cfg.blockDim = dim3(NUM_THREADS);
            cfg.dynamicSmemBytes = SMEM2_BYTES;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = 2;
            at[0].val.clusterDim.y = 1;
            at[0].val.clusterDim.z = 1;
            cfg.attrs = at;
            cfg.numAttrs = 1;
            int c = 0;
            if (cudaOccupancyMaxActiveClusters(&c, gemm_tc2_kernel<false, false>, &cfg) == cudaSuccess) n = c;
            else cudaGetLastError();
            cfg.blockDim = dim3(tc2q_threads(8, 2));  // the QUANT variants (more threads, more shared memory) must fit as often
            cfg.dynamicSmemBytes = 227 * 1024;
            int cq = 0;
            if (cudaOccupancyMaxActiveClusters(&cq, gemm_tc2_kernel<false, true, 5, 2, 8>, &cfg) == cudaSuccess) n = std::min(n, cq);
            else { cudaGetLastError(); n = 0; }
            if (std::getenv("ACE_B200_DEBUG")) std::fprintf(stderr, "[ace_b200] max active CTA pairs: %d\n", n);
        } else cudaGetLastError();
        return cudaSuccess;
    }, &result);
    return result;
}

cudaError_t launch_tc2(const CUtensorMap& tm, const CUtensorMap& tmo, const KParams& kp, int clusters, cudaStream_t stream) {
    cudaLaunchConfig_t cfg = {};
    const bool quant = kp.wtype != W_F16;
    cfg.gridDim = dim3((unsigned)(2 * clusters));
    // QUANT variant: two decoder groups + 4 epilogue warps by default; ACE_B200_GEMM_QDEC=1: one group + 8 epilogue warps (measured
    // decode-bound: gate_up at M = 12 000 1057 vs 1308 TFLOP/s, profiles/r2_quant_sweep_decgroups_i.log).  Ring: the deepest (5, else 4
    // stages) that fits next to the type's packed-code slots
    // Epilogue width: 8 warps (18-warp CTA, 96 registers) where the epilogue is the critical resource — the head-norm epilogue of
    // the Q/K/V projections, and launches of <= 6144 rows where few tiles per cluster leave it exposed — else 4 warps (14-warp CTA,
    // 128 registers).  Measured per shape: profiles/r2_quant_sweep_epi_j.log (QKV +7..16 %, M = 3000 o-proj / down +13..20 %,
    // gate_up at M = 12 000 -5 %).  ACE_B200_GEMM_QEPI=4/8 forces one.
    int dec_groups = 2;
    int epi_warps = (kp.epi.mode == EPI_HEADNORM || (long)kp.B * kp.l_out <= 6144) ? 8 : 4;
    if (const char* e = std::getenv("ACE_B200_GEMM_QDEC")) dec_groups = std::atoi(e) == 1 ? 1 : 2;
    if (const char* e = std::getenv("ACE_B200_GEMM_QEPI")) epi_warps = std::atoi(e) == 8 ? 8 : 4;
    if (dec_groups == 1) epi_warps = 8;
    int ns = quant && smem2q_bytes(5, epi_warps, kp.tile_bytes) <= 227 * 1024 ? 5 : 4;
    if (const char* e = std::getenv("ACE_B200_GEMM_QNS")) {
        if (std::atoi(e) == 4) ns = 4;
        if (std::atoi(e) == 5 && quant && smem2q_bytes(5, epi_warps, kp.tile_bytes) <= 227 * 1024) ns = 5;
    }
    // publish mode (decoded rows reach sB through the bulk-copy engine): two decoder groups, 4-stage ring (a slot must always be
    // decoded by the same group), 32 KB of staging; 8 epilogue warps only where that still fits (not Q8_0's 8.5 KB code slots)
    bool pub = quant && dec_groups == 2 && kp.qpub != 0;
    if (pub) {
        ns = 4;
        if (smem2q_bytes(4, epi_warps, kp.tile_bytes, true) > 227 * 1024) epi_warps = 4;
        if (smem2q_bytes(4, epi_warps, kp.tile_bytes, true) > 227 * 1024) pub = false;
    }
    KParams kq = kp;
    kq.qpub = pub ? 1 : 0;
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = quant ? smem2q_bytes(ns, epi_warps, kp.tile_bytes, pub) : SMEM2_BYTES;
    cfg.stream = stream;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    if (quant) {
        if (dec_groups == 1) return ns == 5 ? launch_tc2q<5, 1, 8>(cfg, tm, tmo, kq) : launch_tc2q<4, 1, 8>(cfg, tm, tmo, kq);
        if (epi_warps == 8) return ns == 5 ? launch_tc2q<5, 2, 8>(cfg, tm, tmo, kq) : launch_tc2q<4, 2, 8>(cfg, tm, tmo, kq);
        return ns == 5 ? launch_tc2q<5, 2, 4>(cfg, tm, tmo, kq) : launch_tc2q<4, 2, 4>(cfg, tm, tmo, kq);
    }
    return kp.warm_epi ? cudaLaunchKernelEx(&cfg, gemm_tc2_kernel<true, false>, tm, tmo, kp) : cudaLaunchKernelEx(&cfg, gemm_tc2_kernel<false, false>, tm, tmo, kp);
}

}  // namespace

static bool make_tmap_sw128(CUtensorMapDataType dt, CUtensorMap* tm, const void* base, int rank, const uint64_t* dims,
                            const uint64_t* strides_bytes, const uint32_t* box);
bool make_tmap_f16_sw128(CUtensorMap* tm, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                         const uint32_t* box) {
    return make_tmap_sw128(CU_TENSOR_MAP_DATA_TYPE_FLOAT16, tm, base, rank, dims, strides_bytes, box);
}
bool make_tmap_f32_sw128(CUtensorMap* tm, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                         const uint32_t* box) {
    return make_tmap_sw128(CU_TENSOR_MAP_DATA_TYPE_FLOAT32, tm, base, rank, dims, strides_bytes, box);
}
static bool make_tmap_sw128(CUtensorMapDataType dt, CUtensorMap* tm, const void* base, int rank, const uint64_t* dims,
                            const uint64_t* strides_bytes, const uint32_t* box) {
    PFN_encodeTiled enc = get_encode_fn();
    if (!enc || rank < 1 || rank > 5) return false;
    cuuint64_t d[5], s[4];
    cuuint32_t bx[5], es[5];
    for (int i = 0; i < rank; ++i) {
        d[i] = dims[i];
        bx[i] = box[i];
        es[i] = 1;
        if (i > 0) s[i - 1] = strides_bytes[i - 1];
    }
    return enc(tm, dt, (cuuint32_t)rank, const_cast<void*>(base), d, s, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

namespace {

template <int BN, bool QUANT, bool WARM>
cudaError_t launch_tc(const CUtensorMap& tm, const CUtensorMap& tmo, const KParams& kp, int grid, cudaStream_t stream) {
    using SL = SmemLayout<BN, QUANT>;
    static DeviceOnce attr;  // one per template instantiation, once per device
    const cudaError_t ea = attr.run([](int&) { return cudaFuncSetAttribute(gemm_tc_kernel<BN, QUANT, WARM>, cudaFuncAttributeMaxDynamicSharedMemorySize, SL::BYTES); });
    if (ea != cudaSuccess) return ea;
    return launch_pdl(gemm_tc_kernel<BN, QUANT, WARM>, dim3((unsigned)grid), dim3(NUM_THREADS), (size_t)SL::BYTES, stream, tm, tmo, kp);
}

cudaError_t launch_tc_slab(const CUtensorMap& tm, const CUtensorMap& tmo, const KParams& kp, int grid, cudaStream_t stream) {
    using SL = SmemLayout<128, false>;
    constexpr int BYTES = 2 * SLAB_BYTES + SL::STAGES * SL::B_STAGE + SL::STG_BYTES + SL::COL_BYTES + 256;
    static_assert(BYTES <= 227 * 1024, "slab layout exceeds shared memory");
    static DeviceOnce attr;
    const cudaError_t ea = attr.run([](int&) { return cudaFuncSetAttribute(gemm_tc_kernel<128, false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, BYTES); });
    if (ea != cudaSuccess) return ea;
    return launch_pdl(gemm_tc_kernel<128, false, false, true>, dim3((unsigned)grid), dim3(NUM_THREADS), (size_t)BYTES, stream, tm, tmo, kp);
}

}  // namespace

int device_sm_count() {
    static DeviceOnce once;
    int sms = 148;
    once.run([](int& v) {
        int dev = 0;
        cudaGetDevice(&dev);
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) v = 148;
        return cudaSuccess;
    }, &sms);
    return sms;
}

bool gemm_force_simt() {
    static int v = -1;
    if (v < 0) {
        const char* e = std::getenv("ACE_B200_GEMM");
        v = (e && std::string(e) == "simt") ? 1 : 0;
    }
    return v == 1;
}

cudaError_t gemm_launch(const GemmProblem& g, cudaStream_t stream, bool simt) {
    const PackedWeight& w = *g.W;
    note_launch();
    // algorithmic HBM bytes: packed weights once + activations once + every output / residual stream the epilogue touches
    const double out_elems = (double)g.B * g.l_out * (g.epi.mode == EPI_SWIGLU ? w.N / 2 : w.N);
    const double out_bytes = out_elems * ((g.epi.out_f32 ? 4.0 : 0.0) + (g.epi.out_f16 ? 2.0 : 0.0) + (g.epi.out_x16 ? 2.0 : 0.0) +
                                          (g.epi.resid ? 4.0 : 0.0) + (g.epi.resid16 ? 2.0 : 0.0));
    ProfScope prof_scope(PROF_GEMM, 2.0 * g.B * (double)g.l_out * w.N * (double)w.K * w.taps,
                         (double)w.bytes + (double)g.B * g.L * g.C * 2.0 + out_bytes, stream);
    KParams kp;
    kp.tma_reduce = 0;
    kp.trace = nullptr;
    kp.dbg = 0;
    if (const char* ed = std::getenv("ACE_B200_GEMM_DBG")) kp.dbg = std::atoi(ed);
    kp.qpub = 0;
    if (const char* ep = std::getenv("ACE_B200_GEMM_QPUB")) kp.qpub = std::atoi(ep) != 0;
    kp.W = static_cast<const uint8_t*>(w.dev);
    kp.wtype = w.wtype;
    kp.tile_bytes = w.tile_bytes;
    if (w.wtype != W_F16 && w.dev_f16 && !simt && !gemm_force_simt() && (long)g.B * g.l_out >= predecode_min_rows()) {
        kp.W = static_cast<const uint8_t*>(w.dev_f16);  // compute-bound: decoded twin through the F16 kernel
        kp.wtype = W_F16;
        kp.tile_bytes = W_SUB_BYTES;
    } else if (w.wtype != W_F16 && !simt && !gemm_force_simt() && decode_pass_min_rows() > 0 && (long)g.B * g.l_out >= decode_pass_min_rows() &&
               w.taps == 1) {
        // decode pass (see decode_pass_min_rows): packed -> F16 image in this stream's scratch, then the F16 kernels
        if (void* sc = decode_scratch(stream, gemm_predecode_bytes(w))) {
            const cudaError_t de = gemm_predecode(w, sc, stream);   // inside this launch's ProfScope: its time counts as GEMM time
            if (de != cudaSuccess) return de;
            kp.W = static_cast<const uint8_t*>(sc);
            kp.wtype = W_F16;
            kp.tile_bytes = W_SUB_BYTES;
        }
    }
    kp.N = w.N;
    kp.n_tiles128 = w.n_tiles;
    kp.kb_per_tap = w.kb_per_tap;
    kp.taps = w.taps;
    kp.kb_total = w.taps * w.kb_per_tap;
    kp.tap_stride = g.tap_stride;
    kp.tap_offset = g.tap_offset;
    kp.B = g.B;
    kp.l_out = g.l_out;
    kp.m_tiles_per_b = (g.l_out + BM - 1) / BM;
    kp.epi = g.epi;
    if (kp.epi.out_valid == 0) kp.epi.out_valid = (long)g.l_out * kp.epi.ld_out;
    {  // specialised store path of the staged epilogue (exact feature set only; anything else runs the generic code)
        const GemmEpilogue& e = kp.epi;
        int v = SV_GENERIC;
        const bool snake = e.snake_alpha != nullptr, plain = e.mode == EPI_STORE && !e.atomic && !e.gate && e.ld_out > 0;
        if (e.resid16 || e.out_x16) {  // 16-bit residual stream (VAE)
            if (plain && snake && e.out_f16 && !e.resid && !e.out_f32)
                v = e.resid16 ? (e.out_x16 ? SV_RES16_X16_F16SNAKE : SV_RES16_F16SNAKE) : SV_X16_F16SNAKE;
        } else if (plain && !e.resid) {
            if (e.out_f16 && !e.out_f32) v = snake ? SV_F16_SNAKE : (e.silu_f16 ? SV_F16_SILU : SV_F16);
            else if (e.out_f32 && !e.out_f16) v = SV_F32;
            else if (e.out_f32 && e.out_f16 && snake) v = SV_F32_F16SNAKE;
        } else if (plain && e.resid && e.out_f16 && snake) {
            v = e.out_f32 ? SV_RES_F32_F16SNAKE : SV_RES_F16SNAKE;
        }
        if (const char* ev = std::getenv("ACE_B200_GEMM_EPIFAST")) if (std::atoi(ev) == 0) v = SV_GENERIC;
        kp.epi_variant = v;
        kp.warm_epi = 1;  // narrowed below once the tile count is known
    }
    // In-place accumulation (out += gate*acc + bias, the DiT residual updates) goes through red.global.add: no
    // read-modify-write latency in the epilogue, and it is what makes a K split legal.
    const bool accumulate = g.epi.mode == EPI_STORE && g.epi.out_f32 && g.epi.resid == g.epi.out_f32 && !g.epi.out_f16;
    const int sms = device_sm_count();
    const int m_tiles = kp.B * kp.m_tiles_per_b;
    // tile shape: 128x256 halves the operand traffic per FLOP (L2 -> SM bandwidth is the limiter of the 128x128 tile);
    // keep 128x128 only where the wider tile would leave most SMs without work
    const long tiles256 = (long)((kp.n_tiles128 + 1) / 2) * m_tiles;
    const long tiles128 = (long)kp.n_tiles128 * m_tiles;
    bool bn256 = kp.n_tiles128 >= 2 && (tiles256 >= sms / 2 || accumulate);
    if (const char* fb = std::getenv("ACE_B200_GEMM_BN")) bn256 = (std::atoi(fb) == 256) && kp.n_tiles128 >= 2;  // tuning knob
    // CTA pair (256 x 256 per cluster): F16 tiles, whole 256-column tiles only; ACE_B200_GEMM_2CTA=0/1 overrides
    const int m_pairs = kp.B * ((g.l_out + 2 * BM - 1) / (2 * BM));
    // Measured (profiles/r1_gemm_sweep_pair2.log): the pair wins once a cluster runs several tiles (M = 6000: +6..17 %),
    // and loses at M = 750 where its longer per-tile ramp (relay hop, cluster syncs) is not amortised -> rows >= 2048.
    // (packed quantised tiles too: gemm_tc2_kernel<., true> decodes them in shared memory; taps == 1 by construction)
    const bool pair_ok = (kp.n_tiles128 & 1) == 0 && kp.n_tiles128 >= 2;
    const long tiles_pair = (long)(kp.n_tiles128 / 2) * m_pairs;
    // ... and below that when every cluster still gets >= ~28 k-blocks of work (M = 750: MLP up-projection N = 12288 792 vs
    // 647 TFLOP/s, down-projection K = 6144 612 vs 526; QKV / o-proj stay single-CTA: profiles/r1_gemm_sweep_issue1.log)
    const int pairs_avail = std::max(1, max_active_pairs());
    bool pair = pair_ok && ((long)g.B * g.l_out >= 2048 || tiles_pair * (long)kp.kb_total >= 28L * pairs_avail);
    // packed quantised tiles: the pair halves the decode work per tensor-core cycle (one sub-tile per CTA per k-block), so it is
    // the route whenever there are two row tiles to pair; one row tile (M <= 128: the HBM-bound GEMVs) stays on the single CTA
    if (kp.wtype != W_F16) pair = pair_ok && (long)g.B * g.l_out > BM;
    if (const char* fp = std::getenv("ACE_B200_GEMM_2CTA")) pair = pair_ok && std::atoi(fp) != 0;
    if (pair && max_active_pairs() < 8) pair = false;
    const int units = pair ? max_active_pairs() : sms;  // schedulable units: clusters or CTAs
    const long tiles_nosplit = pair ? tiles_pair : (bn256 ? tiles256 : tiles128);
    // split-K (g.split_k == 0: automatic): fill the SMs when one wave of tiles does not, >= 8 k-blocks per split
    int sk = g.split_k > 0 ? g.split_k : 1;
    if (g.split_k == 0 && accumulate) {
        sk = (int)((units + tiles_nosplit / 2) / tiles_nosplit);
        sk = std::max(1, std::min(sk, 4));
        while (sk > 1 && kp.kb_total / sk < 8) --sk;
    }
    if (const char* fs = std::getenv("ACE_B200_GEMM_SPLITK")) {  // tuning knob: only honoured where a split is legal
        if (accumulate) sk = std::max(1, std::atoi(fs));
    }
    if (sk > 1 && !accumulate) return cudaErrorInvalidValue;
    kp.split_k = std::min(sk, kp.kb_total);
    {  // bands of row tiles whose activations (~32 MB) stay L2-resident across the weight tiles (tile_coords); ACE_B200_GEMM_BAND_MB=0: off
        double band_mb = 32.0;
        if (const char* eb = std::getenv("ACE_B200_GEMM_BAND_MB")) band_mb = std::atof(eb);
        const int mt_all = pair ? m_pairs : m_tiles;
        const double tile_bytes_a = (double)(pair ? 2 * BM : BM) * g.C * 2.0;
        long bm = band_mb > 0 ? (long)(band_mb * 1048576.0 / tile_bytes_a) : (long)mt_all;
        // (a band may be narrower than a wave: the wave then spans several weight tiles of the same rows, all L2 hits; the weights
        // themselves — <= 50 MB per matrix — stay L2-resident next to the band)
        kp.band_m = (int)std::max<long>(2, std::min<long>(bm, mt_all));
    }
    {  // the dry-run epilogue pass pays off only where a CTA runs a handful of tiles (it costs one extra epilogue per CTA)
        const long work_tiles = tiles_nosplit * kp.split_k;
        kp.warm_epi = work_tiles <= 4L * units ? 1 : 0;
        if (const char* ew = std::getenv("ACE_B200_GEMM_WARM")) kp.warm_epi = std::atoi(ew) != 0;
    }
    // in-place accumulation always goes through red.add, split or not: fire-and-forget, where a read-add-write would put
    // an L2 round trip per row group on the epilogue's critical path (trace: 8.0 k vs 2.2 k cycles per 32-column chunk)
    if (accumulate) kp.epi.atomic = 1;
    if (const char* fa = std::getenv("ACE_B200_GEMM_ATOMIC")) {  // tuning knob: 0 = plain read-add-write when K is not split
        if (std::atoi(fa) == 0 && accumulate && kp.split_k == 1) kp.epi.atomic = 0;
    }
    if (g.C != w.K) return cudaErrorInvalidValue;

    if (simt || gemm_force_simt()) {
        kp.split_k = 1;
        const long total = (long)g.B * g.l_out * ((w.N + 31) / 32);
        const int threads = 128;
        const bool headnorm = kp.epi.mode == EPI_HEADNORM;
        if (headnorm) kp.epi.mode = EPI_STORE;  // raw projection, then the stand-alone norm/RoPE kernel (independent path)
        gemm_simt_kernel<<<(unsigned)((total + threads - 1) / threads), threads, 0, stream>>>(g.A, g.a_lstride, g.a_bstride,
                                                                                            g.L, g.C, kp);
        if (headnorm) {
            const GemmEpilogue& e = g.epi;
            if (e.hn_na > 0) launch_headnorm_rope(e.out_f16, e.ld_out, 0, e.hn_na / 128, e.hn_wa, e.rope, g.B, g.l_out, e.hn_eps, stream);
            if (e.hn_nb > 0) launch_headnorm_rope(e.out_f16, e.ld_out, e.hn_na, e.hn_nb / 128, e.hn_wb, e.rope, g.B, g.l_out, e.hn_eps, stream);
        }
        return cudaGetLastError();
    }

    const long tiles = tiles_nosplit * kp.split_k;
    // activation slab (see gemm_tc_kernel): stride-d taps over <= 128 channels, 128-column tiles, persistent launches only
    kp.slab_rows = 0;
    {
        const int rows = BM + (w.taps - 1) * g.tap_stride;
        bool slab = kp.wtype == W_F16 && !pair && !bn256 && kp.split_k == 1 && w.taps > 1 && g.tap_stride >= 1 && kp.kb_per_tap <= 2 &&
                    rows <= SLAB_MAX_ROWS && !kp.warm_epi && tiles >= 2L * sms;
        if (const char* fs = std::getenv("ACE_B200_GEMM_SLAB")) slab = slab && std::atoi(fs) != 0;
        if (slab) kp.slab_rows = rows;
    }
    CUtensorMap tm;
    if (!make_tmap_a(g, &tm, kp.slab_rows ? kp.slab_rows : BM)) return cudaErrorInvalidValue;
    kp.trace = nullptr;
    static long long* trace_bufs[ACE_MAX_DEVICES] = {};  // debug aid (ACE_B200_GEMM_TRACE), one buffer per device
    int trace_dev = 0;
    cudaGetDevice(&trace_dev);
    long long*& trace_buf = trace_bufs[trace_dev & (ACE_MAX_DEVICES - 1)];
    const bool tracing = std::getenv("ACE_B200_GEMM_TRACE") != nullptr && !pair;
    if (tracing) {
        if (!trace_buf) cudaMalloc(&trace_buf, (size_t)sms * TR_N * sizeof(long long));
        cudaMemsetAsync(trace_buf, 0, (size_t)sms * TR_N * sizeof(long long), stream);
        kp.trace = trace_buf;
    }
    struct TraceDump {
        bool on; long long* buf; int ctas; cudaStream_t st; int M, N, K, bn, sk;
        ~TraceDump() {
            if (!on) return;
            cudaStreamSynchronize(st);
            std::vector<long long> h((size_t)ctas * TR_N);
            cudaMemcpy(h.data(), buf, h.size() * sizeof(long long), cudaMemcpyDeviceToHost);
            double acc[TR_N] = {0};
            double waited = 0;
            for (int c = 0; c < ctas; ++c) {
                for (int i = 0; i < TR_N; ++i) acc[i] += (double)(h[(size_t)c * TR_N + i] - h[(size_t)c * TR_N + TR_START]);
                waited += (double)h[(size_t)c * TR_N + 15];
            }
            std::fprintf(stderr, "[gemm trace] M=%d N=%d K=%d bn=%d sk=%d ctas=%d | cycles from CTA start (mean): setup %.0f first_full %.0f last_mma_issued %.0f "
                                 "acc_ready(last tile) %.0f ld0 %.0f chunk0 %.0f epi_done(last tile) %.0f end %.0f\n", M, N, K, bn, sk, ctas, acc[TR_SETUP] / ctas,
                         acc[TR_FIRST_FULL] / ctas, acc[TR_LAST_MMA] / ctas, acc[TR_ACC_READY] / ctas, acc[TR_EPI_LD0] / ctas, acc[TR_EPI_CHUNK0] / ctas,
                         acc[TR_EPI_DONE] / ctas, acc[TR_END] / ctas);
            if (acc[TR_T8_MMA_DONE] > 0)
                std::fprintf(stderr, "             9th tile of a CTA, cycles from the MMA thread reaching it: accumulator free +%.0f, MMAs issued +%.0f | epilogue warp: "
                                     "starts waiting %+.0f, accumulator ready %+.0f, done %+.0f | MMA thread waited %.0f cycles for operands\n", (acc[TR_T8_MMA_GO] - acc[TR_T8_MMA_WAIT]) / ctas,
                             (acc[TR_T8_MMA_DONE] - acc[TR_T8_MMA_WAIT]) / ctas, (acc[TR_T8_EPI_WAIT] - acc[TR_T8_MMA_WAIT]) / ctas,
                             (acc[TR_T8_EPI_GO] - acc[TR_T8_MMA_WAIT]) / ctas, (acc[TR_T8_EPI_DONE] - acc[TR_T8_MMA_WAIT]) / ctas, waited / ctas);
        }
    } trace_dump{tracing, trace_buf, (int)(tiles < sms ? tiles : sms), stream, g.B * g.l_out, w.N, w.K * w.taps, bn256 ? 256 : 128, kp.split_k};
    // accumulate epilogue as TMA reduce-add tiles (plain [B, l_out, N] fp32 output, no crop shift); ACE_B200_GEMM_TMARED=0 = red.add
    CUtensorMap tmo = tm;
    kp.tma_reduce = 0;
    {
        const char* ft = std::getenv("ACE_B200_GEMM_TMARED");
        const bool want = !(ft && ft[0] == '0');
        const GemmEpilogue& e = kp.epi;
        if (want && e.atomic && e.mode == EPI_STORE && e.out_shift == 0 && e.out_valid == (long)g.l_out * e.ld_out && (e.ld_out & 3) == 0 &&
            (e.out_bstride & 3) == 0 && (reinterpret_cast<uintptr_t>(e.out_f32) & 15) == 0) {
            const uint64_t dims[3] = {(uint64_t)w.N, (uint64_t)g.l_out, (uint64_t)g.B};
            const uint64_t strides[2] = {(uint64_t)e.ld_out * 4, (uint64_t)(g.B > 1 ? e.out_bstride : (long)g.l_out * e.ld_out) * 4};
            const uint32_t box[3] = {32, 32, 1};
            if (make_tmap_f32_sw128(&tmo, e.out_f32, 3, dims, strides, box)) kp.tma_reduce = 1;
        }
    }
    if (pair) return launch_tc2(tm, tmo, kp, (int)(tiles < units ? tiles : units), stream);
    const int grid = (int)(tiles < sms ? tiles : sms);
    const bool quant = kp.wtype != W_F16;
    if (kp.slab_rows) return launch_tc_slab(tm, tmo, kp, grid, stream);
    if (kp.warm_epi) {
        if (bn256) return quant ? launch_tc<256, true, true>(tm, tmo, kp, grid, stream) : launch_tc<256, false, true>(tm, tmo, kp, grid, stream);
        return quant ? launch_tc<128, true, true>(tm, tmo, kp, grid, stream) : launch_tc<128, false, true>(tm, tmo, kp, grid, stream);
    }
    if (bn256) return quant ? launch_tc<256, true, false>(tm, tmo, kp, grid, stream) : launch_tc<256, false, false>(tm, tmo, kp, grid, stream);
    return quant ? launch_tc<128, true, false>(tm, tmo, kp, grid, stream) : launch_tc<128, false, false>(tm, tmo, kp, grid, stream);
}

}  // namespace ace
