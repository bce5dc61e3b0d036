// Host side of the weight formats: ggml block decoders, Q8_0 quantiser and the tile packer (see wpack.h).
#include "wpack.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <thread>

namespace ace {

// ------------------------------------------------------------------ scalar conversions
float f16_to_f32(uint16_t h) {
    const uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1Fu;
    uint32_t man = h & 0x3FFu;
    uint32_t bits;
    if (exp == 0) {
        if (man == 0) {
            bits = sign;
        } else {  // subnormal: renormalise
            int e = -1;
            do {
                man <<= 1;
                ++e;
            } while ((man & 0x400u) == 0);
            man &= 0x3FFu;
            bits = sign | ((uint32_t)(127 - 15 - e) << 23) | (man << 13);
        }
    } else if (exp == 31) {
        bits = sign | 0x7F800000u | (man << 13);
    } else {
        bits = sign | ((exp + 112u) << 23) | (man << 13);
    }
    float f;
    std::memcpy(&f, &bits, 4);
    return f;
}

uint16_t f32_to_f16(float f) {
    uint32_t x;
    std::memcpy(&x, &f, 4);
    const uint16_t sign = (uint16_t)((x >> 16) & 0x8000u);
    const uint32_t absx = x & 0x7FFFFFFFu;
    if (absx >= 0x7F800000u) {  // inf / nan
        return (uint16_t)(sign | 0x7C00u | ((absx > 0x7F800000u) ? 0x200u : 0u));
    }
    if (absx >= 0x477FF000u) {  // rounds to >= 65520 -> inf
        return (uint16_t)(sign | 0x7C00u);
    }
    if (absx < 0x33000001u) {  // < 2^-25 (or exactly 2^-25 which ties to even = 0)
        return sign;
    }
    int32_t exp = (int32_t)(absx >> 23) - 127;
    uint32_t man = (absx & 0x7FFFFFu) | 0x800000u;
    int shift;
    uint32_t base;
    if (exp < -14) {  // subnormal half
        shift = 13 + (-14 - exp);
        base = 0;
    } else {
        shift = 13;
        base = ((uint32_t)(exp + 15) << 10) - 0x400u;  // implicit bit adds the 0x400 back
    }
    uint32_t q = man >> shift;
    const uint32_t rem = man & ((1u << shift) - 1u);
    const uint32_t half = 1u << (shift - 1);
    if (rem > half || (rem == half && (q & 1u))) {
        ++q;
    }
    return (uint16_t)(sign | (base + q));
}

float bf16_to_f32(uint16_t h) {
    const uint32_t bits = (uint32_t)h << 16;
    float f;
    std::memcpy(&f, &bits, 4);
    return f;
}

// ------------------------------------------------------------------ ggml block geometry
int ggml_block_elems(int t) {
    switch (t) {
        case GGML_F32: case GGML_F16: case GGML_BF16: return 1;
        case GGML_Q8_0: return 32;
        case GGML_Q4_K: case GGML_Q6_K: return 256;
    }
    return 0;
}

static int ggml_block_bytes(int t) {
    switch (t) {
        case GGML_F32: return 4;
        case GGML_F16: case GGML_BF16: return 2;
        case GGML_Q8_0: return 34;
        case GGML_Q4_K: return 144;
        case GGML_Q6_K: return 210;
    }
    return 0;
}

size_t ggml_row_bytes(int t, int64_t n) {
    const int be = ggml_block_elems(t);
    if (be == 0 || n % be != 0) return 0;
    return (size_t)(n / be) * (size_t)ggml_block_bytes(t);
}

const char* ggml_type_name(int t) {
    switch (t) {
        case GGML_F32: return "F32";
        case GGML_F16: return "F16";
        case GGML_BF16: return "BF16";
        case GGML_Q8_0: return "Q8_0";
        case GGML_Q4_K: return "Q4_K";
        case GGML_Q6_K: return "Q6_K";
    }
    return "?";
}

static inline uint16_t rd16(const uint8_t* p) { return (uint16_t)(p[0] | (p[1] << 8)); }

// 6-bit (scale, min) pair j of a q4_K super-block (ggml-quants.c get_scale_min_k4)
static inline void q4k_scale_min(int j, const uint8_t* q, uint8_t& sc, uint8_t& m) {
    if (j < 4) {
        sc = q[j] & 63;
        m = q[j + 4] & 63;
    } else {
        sc = (uint8_t)((q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4));
        m = (uint8_t)((q[j + 4] >> 4) | ((q[j] >> 6) << 4));
    }
}

// Generic "codes + group scales" view of one row; every quantised type decodes as
//   w[k] = scale[k / group] * code[k] - offs[k / group]
struct RowCodes {
    std::vector<int16_t> code;
    std::vector<float> scale, offs;
    int group = 32;
};

static bool decode_row_codes(int t, const uint8_t* src, int64_t n, RowCodes& rc) {
    rc.code.resize((size_t)n);
    if (t == GGML_Q8_0) {
        rc.group = 32;
        rc.scale.assign((size_t)(n / 32), 0.f);
        rc.offs.assign((size_t)(n / 32), 0.f);
        for (int64_t b = 0; b < n / 32; ++b) {
            const uint8_t* blk = src + b * 34;
            rc.scale[(size_t)b] = f16_to_f32(rd16(blk));
            for (int i = 0; i < 32; ++i) rc.code[(size_t)(b * 32 + i)] = (int8_t)blk[2 + i];
        }
        return true;
    }
    if (t == GGML_Q4_K) {
        rc.group = 32;
        rc.scale.assign((size_t)(n / 32), 0.f);
        rc.offs.assign((size_t)(n / 32), 0.f);
        for (int64_t b = 0; b < n / 256; ++b) {
            const uint8_t* blk = src + b * 144;
            const float d = f16_to_f32(rd16(blk));
            const float dmin = f16_to_f32(rd16(blk + 2));
            const uint8_t* scales = blk + 4;
            const uint8_t* q = blk + 16;
            for (int j = 0; j < 4; ++j) {  // 64 weights per 32 bytes: low nibbles then high nibbles
                uint8_t sc, m;
                q4k_scale_min(2 * j, scales, sc, m);
                rc.scale[(size_t)(b * 8 + 2 * j)] = d * sc;
                rc.offs[(size_t)(b * 8 + 2 * j)] = dmin * m;
                q4k_scale_min(2 * j + 1, scales, sc, m);
                rc.scale[(size_t)(b * 8 + 2 * j + 1)] = d * sc;
                rc.offs[(size_t)(b * 8 + 2 * j + 1)] = dmin * m;
                for (int l = 0; l < 32; ++l) {
                    rc.code[(size_t)(b * 256 + j * 64 + l)] = q[j * 32 + l] & 0xF;
                    rc.code[(size_t)(b * 256 + j * 64 + 32 + l)] = q[j * 32 + l] >> 4;
                }
            }
        }
        return true;
    }
    if (t == GGML_Q6_K) {
        rc.group = 16;
        rc.scale.assign((size_t)(n / 16), 0.f);
        rc.offs.assign((size_t)(n / 16), 0.f);
        for (int64_t b = 0; b < n / 256; ++b) {
            const uint8_t* blk = src + b * 210;
            const uint8_t* ql = blk;
            const uint8_t* qh = blk + 128;
            const int8_t* sc = reinterpret_cast<const int8_t*>(blk + 192);
            const float d = f16_to_f32(rd16(blk + 208));
            for (int s = 0; s < 16; ++s) rc.scale[(size_t)(b * 16 + s)] = d * sc[s];
            for (int h = 0; h < 2; ++h) {  // two halves of 128 weights
                for (int l = 0; l < 32; ++l) {
                    const uint8_t hb = qh[h * 32 + l];
                    const int base = (int)(b * 256 + h * 128);
                    rc.code[(size_t)(base + l)] = (int16_t)(((ql[h * 64 + l] & 0xF) | (((hb >> 0) & 3) << 4)) - 32);
                    rc.code[(size_t)(base + 32 + l)] = (int16_t)(((ql[h * 64 + 32 + l] & 0xF) | (((hb >> 2) & 3) << 4)) - 32);
                    rc.code[(size_t)(base + 64 + l)] = (int16_t)(((ql[h * 64 + l] >> 4) | (((hb >> 4) & 3) << 4)) - 32);
                    rc.code[(size_t)(base + 96 + l)] = (int16_t)(((ql[h * 64 + 32 + l] >> 4) | (((hb >> 6) & 3) << 4)) - 32);
                }
            }
        }
        return true;
    }
    return false;
}

bool dequant_row(int t, const void* src_v, float* dst, int64_t n) {
    const uint8_t* src = static_cast<const uint8_t*>(src_v);
    switch (t) {
        case GGML_F32:
            std::memcpy(dst, src, (size_t)n * 4);
            return true;
        case GGML_F16:
            for (int64_t i = 0; i < n; ++i) dst[i] = f16_to_f32(rd16(src + 2 * i));
            return true;
        case GGML_BF16:
            for (int64_t i = 0; i < n; ++i) dst[i] = bf16_to_f32(rd16(src + 2 * i));
            return true;
        case GGML_Q8_0:
        case GGML_Q4_K:
        case GGML_Q6_K: {
            if (n % ggml_block_elems(t) != 0) return false;
            RowCodes rc;
            if (!decode_row_codes(t, src, n, rc)) return false;
            for (int64_t i = 0; i < n; ++i) {
                const size_t g = (size_t)(i / rc.group);
                // same operation order as ggml: d*q (q8_0), d1*q - m1 (q4_K), d*sc*q (q6_K)
                dst[i] = rc.scale[g] * (float)rc.code[(size_t)i] - rc.offs[g];
            }
            return true;
        }
    }
    return false;
}

void quantize_row_q8_0(const float* x, void* dst_v, int64_t n) {
    uint8_t* dst = static_cast<uint8_t*>(dst_v);
    for (int64_t b = 0; b < n / 32; ++b) {
        float amax = 0.f;
        for (int i = 0; i < 32; ++i) amax = std::fmax(amax, std::fabs(x[b * 32 + i]));
        const float d = amax / 127.0f;
        const float id = d != 0.f ? 1.0f / d : 0.0f;
        const uint16_t dh = f32_to_f16(d);
        uint8_t* blk = dst + b * 34;
        blk[0] = (uint8_t)(dh & 0xFF);
        blk[1] = (uint8_t)(dh >> 8);
        for (int i = 0; i < 32; ++i) blk[2 + i] = (uint8_t)(int8_t)std::lroundf(x[b * 32 + i] * id);
    }
}

// ------------------------------------------------------------------ tile packer
// n-tiles are independent: pack them on all host cores (load time of a 1.6 G-parameter model is dominated by this)
template <typename F>
static void parallel_for(int n, F&& fn) {
    unsigned hw = std::thread::hardware_concurrency();
    if (hw == 0) hw = 4;
    const int nthreads = (int)std::min<unsigned>(hw, (unsigned)std::max(1, n));
    if (nthreads <= 1) {
        for (int i = 0; i < n; ++i) fn(i);
        return;
    }
    std::atomic<int> next{0};
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t)
        th.emplace_back([&] {
            for (int i = next.fetch_add(1); i < n; i = next.fetch_add(1)) fn(i);
        });
    for (auto& t : th) t.join();
}

static inline void put16(uint8_t* p, uint16_t v) {
    p[0] = (uint8_t)(v & 0xFF);
    p[1] = (uint8_t)(v >> 8);
}

static void init_meta(PackedWeight& m, int wtype, int N, int taps, int K) {
    m.dev = nullptr;
    m.wtype = wtype;
    m.N = N;
    m.K = K;
    m.taps = taps;
    m.kb_per_tap = (K + TILE_K - 1) / TILE_K;
    m.n_tiles = (N + TILE_N - 1) / TILE_N;
    m.tile_bytes = wtype_tile_bytes(wtype);
    m.bytes = (size_t)m.n_tiles * m.taps * m.kb_per_tap * m.tile_bytes;
}

// F16 tile: UMMA K-major SWIZZLE_128B image. element (r, k) at r*128 + ((k/8) ^ (r&7))*16 + (k%8)*2
static void pack_f16_tile(uint8_t* tile, const float* w, long row_stride, int rows, int cols) {
    std::memset(tile, 0, 16384);
    for (int r = 0; r < rows; ++r) {
        for (int k = 0; k < cols; ++k) {
            const size_t off = (size_t)r * 128 + (size_t)(((k >> 3) ^ (r & 7)) << 4) + (size_t)(k & 7) * 2;
            put16(tile + off, f32_to_f16(w[(size_t)r * row_stride + k]));
        }
    }
}

// pack the 8 codes of columns k0..k0+7 into one nibble word so that the kernel's
// ((w >> 4i) & 0x000F000F) yields the pair (k0+2i, k0+2i+1): nibble i <- code 2i, nibble i+4 <- code 2i+1
static inline uint32_t nibble_word(const int* c) {
    uint32_t w = 0;
    for (int i = 0; i < 4; ++i) {
        w |= (uint32_t)(c[2 * i] & 0xF) << (4 * i);
        w |= (uint32_t)(c[2 * i + 1] & 0xF) << (4 * (i + 4));
    }
    return w;
}
static inline void put32(uint8_t* p, uint32_t v) {
    p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24);
}

static void pack_quant_tile(uint8_t* tile, int wtype, const std::vector<RowCodes>& rows, int r0, int nrows, int k0) {
    std::memset(tile, 0, (size_t)wtype_tile_bytes(wtype));
    for (int r = 0; r < nrows; ++r) {
        const RowCodes& rc = rows[(size_t)(r0 + r)];
        if (wtype == W_Q8_0) {
            for (int j = 0; j < 4; ++j)
                for (int i = 0; i < 16; ++i)
                    tile[(size_t)j * 2048 + (size_t)r * 16 + i] = (uint8_t)(int8_t)rc.code[(size_t)(k0 + 16 * j + i)];
            for (int g = 0; g < 2; ++g)
                put16(tile + 8192 + (size_t)r * 4 + g * 2, f32_to_f16(rc.scale[(size_t)(k0 / 32 + g)]));
        } else if (wtype == W_Q4_K) {
            for (int j = 0; j < 2; ++j) {
                for (int wi = 0; wi < 4; ++wi) {
                    int c[8];
                    for (int i = 0; i < 8; ++i) c[i] = rc.code[(size_t)(k0 + 32 * j + 8 * wi + i)];
                    put32(tile + (size_t)j * 2048 + (size_t)r * 16 + wi * 4, nibble_word(c));
                }
                put16(tile + 4096 + (size_t)r * 8 + j * 4, f32_to_f16(rc.scale[(size_t)(k0 / 32 + j)]));
                put16(tile + 4096 + (size_t)r * 8 + j * 4 + 2, f32_to_f16(rc.offs[(size_t)(k0 / 32 + j)]));
            }
        } else {  // W_Q6_K: code+32 in 0..63 -> low nibble plane + 2-bit plane
            uint32_t hi[4] = {0, 0, 0, 0};
            for (int g = 0; g < 8; ++g) {  // word g covers columns 8g..8g+7
                int lo[8];
                for (int i = 0; i < 8; ++i) {
                    const int u = rc.code[(size_t)(k0 + 8 * g + i)] + 32;
                    lo[i] = u & 0xF;
                    const uint32_t h2 = (uint32_t)(u >> 4) & 3u;
                    // pair index p = i/2, element e = i%2 (0 -> low half, 1 -> high half of the half2)
                    const int p = i >> 1, e = i & 1;
                    const int bit = (g & 1 ? 8 : 0) + (e ? 16 : 0) + 2 * p;
                    hi[g >> 1] |= h2 << bit;
                }
                put32(tile + (size_t)(g >> 2) * 2048 + (size_t)r * 16 + (size_t)(g & 3) * 4, nibble_word(lo));
            }
            for (int q = 0; q < 4; ++q) {
                put32(tile + 4096 + (size_t)r * 16 + q * 4, hi[q]);
                put16(tile + 6144 + (size_t)r * 8 + q * 2, f32_to_f16(rc.scale[(size_t)(k0 / 16 + q)]));
            }
        }
    }
}

bool pack_tiles_ggml(const void* rows_v, int ggml_type, int N, int K, std::vector<uint8_t>& out, PackedWeight& meta,
                     std::string& err) {
    int wtype;
    switch (ggml_type) {
        case GGML_Q8_0: wtype = W_Q8_0; break;
        case GGML_Q4_K: wtype = W_Q4_K; break;
        case GGML_Q6_K: wtype = W_Q6_K; break;
        default: err = std::string("pack_tiles_ggml: not a quantised type: ") + ggml_type_name(ggml_type); return false;
    }
    if (K % TILE_K != 0 || K % ggml_block_elems(ggml_type) != 0) {
        err = "pack_tiles_ggml: K not a multiple of the tile / block size";
        return false;
    }
    const size_t rb = ggml_row_bytes(ggml_type, K);
    const uint8_t* rows = static_cast<const uint8_t*>(rows_v);
    init_meta(meta, wtype, N, 1, K);
    out.assign(meta.bytes, 0);
    parallel_for(meta.n_tiles, [&](int nt) {
        std::vector<RowCodes> rc((size_t)TILE_N);
        const int r0 = nt * TILE_N;
        const int nr = std::min(TILE_N, N - r0);
        for (int r = 0; r < nr; ++r) decode_row_codes(ggml_type, rows + (size_t)(r0 + r) * rb, K, rc[(size_t)r]);
        for (int kb = 0; kb < meta.kb_per_tap; ++kb)
            pack_quant_tile(out.data() + meta.tile_offset(nt, kb), wtype, rc, 0, nr, kb * TILE_K);
    });
    return true;
}

// ------------------------------------------------------------------ K-quant row quantisers (on-line ACE_GGML_*_WEIGHT_QTYPE)
// Restatement of ggml's reference quantisers — quantize_row_q4_K_ref / quantize_row_q6_K_ref with make_qkx2_quants /
// make_qx_quants (third_party/ggml/src/ggml-quants.c:444-518, 622-701, 1280-1350, 1692-1760) — with the float operations in
// the same order, so the blocks are bit-identical to what ggml_quantize_chunk writes (tests/test_wpack_cpu.py checks that
// against the reference library).  The reference runs them at load when the env switch names a K-quant
// (acestep_dit_model.cpp:156-192); this library does the same.
namespace {

inline int round_magic(float v) {  // ggml's nearest_int: round-to-nearest-even through the 1.5 * 2^23 trick
    const float t = v + 12582912.f;
    int i;
    std::memcpy(&i, &t, sizeof(i));
    return (i & 0x007fffff) - 0x00400000;
}

// asymmetric 4-bit group fit: codes in [0, nmax], returns the scale, *neg_min = -offset (weighted least squares over a
// small grid of inverse scales)
float fit_scale_min(int n, int nmax, const float* x, const float* wgt, uint8_t* L, float* neg_min, uint8_t* trial, float rmin,
                    float rdelta, int nstep) {
    float lo = x[0], hi = x[0], sum_w = wgt[0], sum_x = sum_w * x[0];
    for (int i = 1; i < n; ++i) {
        if (x[i] < lo) lo = x[i];
        if (x[i] > hi) hi = x[i];
        const float w = wgt[i];
        sum_w += w;
        sum_x += w * x[i];
    }
    if (lo > 0) lo = 0;
    if (hi == lo) {
        for (int i = 0; i < n; ++i) L[i] = 0;
        *neg_min = -lo;
        return 0.f;
    }
    float iscale = nmax / (hi - lo);
    float scale = 1 / iscale;
    float best = 0;
    for (int i = 0; i < n; ++i) {
        const int l = round_magic(iscale * (x[i] - lo));
        L[i] = (uint8_t)std::max(0, std::min(nmax, l));
        float diff = scale * L[i] + lo - x[i];
        diff = diff * diff;
        best += wgt[i] * diff;
    }
    for (int is = 0; is <= nstep; ++is) {
        iscale = (rmin + rdelta * is + nmax) / (hi - lo);
        float sl = 0, sl2 = 0, sxl = 0;
        for (int i = 0; i < n; ++i) {
            int l = round_magic(iscale * (x[i] - lo));
            l = std::max(0, std::min(nmax, l));
            trial[i] = (uint8_t)l;
            const float w = wgt[i];
            sl += w * l;
            sl2 += w * l * l;
            sxl += w * l * x[i];
        }
        const float D = sum_w * sl2 - sl * sl;
        if (D > 0) {
            float sc = (sum_w * sxl - sum_x * sl) / D;
            float mn = (sl2 * sum_x - sl * sxl) / D;
            if (mn > 0) {
                mn = 0;
                sc = sxl / sl2;
            }
            float cur = 0;
            for (int i = 0; i < n; ++i) {
                float diff = sc * trial[i] + mn - x[i];
                diff = diff * diff;
                cur += wgt[i] * diff;
            }
            if (cur < best) {
                for (int i = 0; i < n; ++i) L[i] = trial[i];
                best = cur;
                scale = sc;
                lo = mn;
            }
        }
    }
    *neg_min = -lo;
    return scale;
}

// symmetric group fit used by Q6_K (x^2 weights, 18 trial scales around -nmax / max)
float fit_scale_sym(int n, int nmax, const float* x, int8_t* L) {
    float mx = 0, amax = 0;
    for (int i = 0; i < n; ++i) {
        const float ax = std::fabs(x[i]);
        if (ax > amax) {
            amax = ax;
            mx = x[i];
        }
    }
    if (amax < 1e-15f) {
        for (int i = 0; i < n; ++i) L[i] = 0;
        return 0.f;
    }
    float iscale = -nmax / mx;
    float sumlx = 0, suml2 = 0;
    for (int i = 0; i < n; ++i) {
        int l = round_magic(iscale * x[i]);
        l = std::max(-nmax, std::min(nmax - 1, l));
        L[i] = (int8_t)(l + nmax);
        const float w = x[i] * x[i];
        sumlx += w * x[i] * l;
        suml2 += w * l * l;
    }
    float scale = suml2 ? sumlx / suml2 : 0.0f;
    float best = scale * sumlx;
    for (int is = -9; is <= 9; ++is) {
        if (is == 0) continue;
        iscale = -(nmax + 0.1f * is) / mx;
        sumlx = suml2 = 0;
        for (int i = 0; i < n; ++i) {
            int l = round_magic(iscale * x[i]);
            l = std::max(-nmax, std::min(nmax - 1, l));
            const float w = x[i] * x[i];
            sumlx += w * x[i] * l;
            suml2 += w * l * l;
        }
        if (suml2 > 0 && sumlx * sumlx > best * suml2) {
            for (int i = 0; i < n; ++i) {
                const int l = round_magic(iscale * x[i]);
                L[i] = (int8_t)(nmax + std::max(-nmax, std::min(nmax - 1, l)));
            }
            scale = sumlx / suml2;
            best = scale * sumlx;
        }
    }
    return scale;
}

}  // namespace

// block_q4_K: fp16 d, fp16 dmin, 12 bytes of packed 6-bit (scale, min) pairs, 128 bytes of nibbles  (ggml-common.h:291-305)
void quantize_row_q4_k(const float* x, void* dst, int64_t n) {
    uint8_t* out = static_cast<uint8_t*>(dst);
    uint8_t L[256], trial[32];
    float wgt[32], mins[8], scales[8];
    for (int64_t b = 0; b < n / 256; ++b, x += 256, out += 144) {
        float max_scale = 0, max_min = 0;
        for (int j = 0; j < 8; ++j) {
            float sum_x2 = 0;
            for (int l = 0; l < 32; ++l) sum_x2 += x[32 * j + l] * x[32 * j + l];
            const float av = std::sqrt(sum_x2 / 32);
            for (int l = 0; l < 32; ++l) wgt[l] = av + std::fabs(x[32 * j + l]);
            scales[j] = fit_scale_min(32, 15, x + 32 * j, wgt, L + 32 * j, &mins[j], trial, -1.f, 0.1f, 20);
            if (scales[j] > max_scale) max_scale = scales[j];
            if (mins[j] > max_min) max_min = mins[j];
        }
        const float inv_scale = max_scale > 0 ? 63.f / max_scale : 0.f;
        const float inv_min = max_min > 0 ? 63.f / max_min : 0.f;
        uint8_t* sc = out + 4;
        std::memset(sc, 0, 12);
        for (int j = 0; j < 8; ++j) {
            uint8_t ls = (uint8_t)round_magic(inv_scale * scales[j]);
            uint8_t lm = (uint8_t)round_magic(inv_min * mins[j]);
            ls = std::min<uint8_t>(63, ls);
            lm = std::min<uint8_t>(63, lm);
            if (j < 4) {
                sc[j] = ls;
                sc[j + 4] = lm;
            } else {
                sc[j + 4] = (uint8_t)((ls & 0xF) | ((lm & 0xF) << 4));
                sc[j - 4] |= (uint8_t)((ls >> 4) << 6);
                sc[j] |= (uint8_t)((lm >> 4) << 6);
            }
        }
        const uint16_t dh = f32_to_f16(max_scale / 63.f), mh = f32_to_f16(max_min / 63.f);
        std::memcpy(out, &dh, 2);
        std::memcpy(out + 2, &mh, 2);
        const float d_all = f16_to_f32(dh), m_all = f16_to_f32(mh);
        for (int j = 0; j < 8; ++j) {
            uint8_t s6, m6;
            q4k_scale_min(j, sc, s6, m6);
            const float d = d_all * s6;
            if (!d) continue;
            const float dm = m_all * m6;
            for (int i = 0; i < 32; ++i) {
                const int l = round_magic((x[32 * j + i] + dm) / d);
                L[32 * j + i] = (uint8_t)std::max(0, std::min(15, l));
            }
        }
        uint8_t* q = out + 16;
        for (int j = 0; j < 256; j += 64, q += 32)
            for (int l = 0; l < 32; ++l) q[l] = (uint8_t)(L[j + l] | (L[j + l + 32] << 4));
    }
}

// block_q6_K: 128 bytes low nibbles, 64 bytes high 2-bit planes, 16 int8 scales, fp16 d  (ggml-common.h:326-335)
void quantize_row_q6_k(const float* x, void* dst, int64_t n) {
    uint8_t* out = static_cast<uint8_t*>(dst);
    int8_t L[256];
    float scales[16];
    for (int64_t b = 0; b < n / 256; ++b, x += 256, out += 210) {
        float max_scale = 0, max_abs = 0;
        for (int ib = 0; ib < 16; ++ib) {
            const float sc = fit_scale_sym(16, 32, x + 16 * ib, L + 16 * ib);
            scales[ib] = sc;
            const float a = std::fabs(sc);
            if (a > max_abs) {
                max_abs = a;
                max_scale = sc;
            }
        }
        if (max_abs < 1e-15f) {
            std::memset(out, 0, 210);
            continue;
        }
        const float iscale = -128.f / max_scale;
        const uint16_t dh = f32_to_f16(1 / iscale);
        std::memcpy(out + 208, &dh, 2);
        int8_t* sc8 = reinterpret_cast<int8_t*>(out + 192);
        for (int ib = 0; ib < 16; ++ib) sc8[ib] = (int8_t)std::min(127, round_magic(iscale * scales[ib]));
        const float d_all = f16_to_f32(dh);
        for (int j = 0; j < 16; ++j) {
            const float d = d_all * sc8[j];
            if (!d) continue;
            for (int i = 0; i < 16; ++i) {
                int l = round_magic(x[16 * j + i] / d);
                l = std::max(-32, std::min(31, l));
                L[16 * j + i] = (int8_t)(l + 32);
            }
        }
        uint8_t* ql = out;
        uint8_t* qh = out + 128;
        for (int j = 0; j < 256; j += 128, ql += 64, qh += 32) {
            for (int l = 0; l < 32; ++l) {
                const uint8_t q1 = L[j + l] & 0xF, q2 = L[j + l + 32] & 0xF, q3 = L[j + l + 64] & 0xF, q4 = L[j + l + 96] & 0xF;
                ql[l] = (uint8_t)(q1 | (q3 << 4));
                ql[l + 32] = (uint8_t)(q2 | (q4 << 4));
                qh[l] = (uint8_t)((L[j + l] >> 4) | ((L[j + l + 32] >> 4) << 2) | ((L[j + l + 64] >> 4) << 4) | ((L[j + l + 96] >> 4) << 6));
            }
        }
    }
}


bool pack_tiles_f32(const float* w, int N, int taps, int K, int wtype, std::vector<uint8_t>& out, PackedWeight& meta,
                    std::string& err) {
    if (wtype == W_Q8_0 || wtype == W_Q4_K || wtype == W_Q6_K) {
        const int blk_elems = wtype == W_Q8_0 ? TILE_K : 256;
        if (taps != 1 || K % blk_elems != 0) {
            err = "pack_tiles_f32: on-line quantisation needs taps == 1 and K % 64 == 0 (K % 256 for the K-quants)";
            return false;
        }
        const int gt = wtype == W_Q8_0 ? GGML_Q8_0 : (wtype == W_Q4_K ? GGML_Q4_K : GGML_Q6_K);
        const size_t rb = ggml_row_bytes(gt, K);
        std::vector<uint8_t> q((size_t)N * rb);
        parallel_for((N + 15) / 16, [&](int blk) {
            for (int r = blk * 16; r < std::min(N, blk * 16 + 16); ++r) {
                const float* src = w + (size_t)r * K;
                uint8_t* dst = q.data() + (size_t)r * rb;
                if (wtype == W_Q8_0) quantize_row_q8_0(src, dst, K);
                else if (wtype == W_Q4_K) quantize_row_q4_k(src, dst, K);
                else quantize_row_q6_k(src, dst, K);
            }
        });
        return pack_tiles_ggml(q.data(), gt, N, K, out, meta, err);
    }
    if (wtype != W_F16) {
        err = "pack_tiles_f32: unknown weight type";
        return false;
    }
    init_meta(meta, W_F16, N, taps, K);
    out.assign(meta.bytes, 0);
    const long row_stride = (long)taps * K;
    parallel_for(meta.n_tiles, [&](int nt) {
        const int r0 = nt * TILE_N;
        const int nr = std::min(TILE_N, N - r0);
        for (int tap = 0; tap < taps; ++tap) {
            for (int kb = 0; kb < meta.kb_per_tap; ++kb) {
                const int k0 = kb * TILE_K;
                const int nc = std::min(TILE_K, K - k0);
                pack_f16_tile(out.data() + meta.tile_offset(nt, tap * meta.kb_per_tap + kb),
                              w + (size_t)r0 * row_stride + (size_t)tap * K + k0, row_stride, nr, nc);
            }
        }
    });
    return true;
}

// ------------------------------------------------------------------ host emulation of the kernel decoder
void unpack_tiles_host(const std::vector<uint8_t>& img, const PackedWeight& m, std::vector<float>& w) {
    const long row_stride = (long)m.taps * m.K;
    w.assign((size_t)m.N * row_stride, 0.f);
    auto rd32 = [](const uint8_t* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); };
    for (int nt = 0; nt < m.n_tiles; ++nt) {
        for (int kbt = 0; kbt < m.taps * m.kb_per_tap; ++kbt) {
            const uint8_t* tile = img.data() + m.tile_offset(nt, kbt);
            const int tap = kbt / m.kb_per_tap, kb = kbt % m.kb_per_tap;
            for (int r = 0; r < TILE_N; ++r) {
                const int n = nt * TILE_N + r;
                if (n >= m.N) break;
                float v[64];
                if (m.wtype == W_F16) {
                    for (int k = 0; k < 64; ++k)
                        v[k] = f16_to_f32(rd16(tile + (size_t)r * 128 + (size_t)(((k >> 3) ^ (r & 7)) << 4) + (size_t)(k & 7) * 2));
                } else if (m.wtype == W_Q8_0) {
                    for (int j = 0; j < 4; ++j) {
                        const float sc = f16_to_f32(rd16(tile + 8192 + (size_t)r * 4 + (j >> 1) * 2));
                        for (int i = 0; i < 16; ++i) {
                            const uint32_t u = (uint8_t)(tile[(size_t)j * 2048 + (size_t)r * 16 + i] ^ 0x80);  // kernel: 1024+u-1152
                            v[16 * j + i] = ((float)u - 128.f) * sc;
                        }
                    }
                } else if (m.wtype == W_Q4_K) {
                    for (int j = 0; j < 2; ++j) {
                        const float ds = f16_to_f32(rd16(tile + 4096 + (size_t)r * 8 + j * 4));
                        const float dm = f16_to_f32(rd16(tile + 4096 + (size_t)r * 8 + j * 4 + 2));
                        for (int wi = 0; wi < 4; ++wi) {
                            const uint32_t wd = rd32(tile + (size_t)j * 2048 + (size_t)r * 16 + wi * 4);
                            for (int i = 0; i < 4; ++i) {
                                const uint32_t t = (wd >> (4 * i)) & 0x000F000Fu;
                                v[32 * j + 8 * wi + 2 * i] = (float)(t & 0xFFFF) * ds - dm;
                                v[32 * j + 8 * wi + 2 * i + 1] = (float)(t >> 16) * ds - dm;
                            }
                        }
                    }
                } else {
                    for (int g = 0; g < 8; ++g) {
                        const uint32_t wd = rd32(tile + (size_t)(g >> 2) * 2048 + (size_t)r * 16 + (size_t)(g & 3) * 4);
                        const uint32_t P = rd32(tile + 4096 + (size_t)r * 16 + (size_t)(g >> 1) * 4);
                        const uint32_t Hs = (g & 1) ? (P >> 4) : (P << 4);
                        const float sc = f16_to_f32(rd16(tile + 6144 + (size_t)r * 8 + (size_t)(g >> 1) * 2));
                        for (int i = 0; i < 4; ++i) {
                            const uint32_t t = ((wd >> (4 * i)) & 0x000F000Fu) | ((Hs >> (2 * i)) & 0x00300030u);
                            v[8 * g + 2 * i] = ((float)(t & 0xFFFF) - 32.f) * sc;
                            v[8 * g + 2 * i + 1] = ((float)(t >> 16) - 32.f) * sc;
                        }
                    }
                }
                for (int k = 0; k < 64; ++k) {
                    const int kk = kb * 64 + k;
                    if (kk < m.K) w[(size_t)n * row_stride + (size_t)tap * m.K + kk] = v[k];
                }
            }
        }
    }
}

}  // namespace ace
