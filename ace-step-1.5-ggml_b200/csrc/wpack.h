// Weight formats: ggml block decoding on the host (file side) and the tile-major packed layout in HBM
// (device side).
//
// File side (what the loaders must accept, bit-exact): F32 / F16 / BF16 / Q8_0 / Q4_K / Q6_K blocks along
// the input dimension   reference: third_party/ggml/src/ggml-common.h:219-223 (q8_0), 291-305 (q4_K),
// 326-335 (q6_K); dequantisers third_party/ggml/src/ggml-quants.c:401-416, 1352-1375, 1762-1792.
//
// Device side: the weight matrix W[N, taps*K] is cut into tiles of 128 rows x 64 columns (one GEMM
// pipeline stage) and every tile is stored as ONE contiguous byte range, laid out in the order the
// kernel consumes it, so a single cp.async.bulk (TMA) brings a tile into shared memory:
//   F16   16384 B  the UMMA K-major / 128B-swizzle image itself (no conversion in the kernel)
//   Q8_0   8704 B  int8 codes [4][128][16] + fp16 d      [128][2]      (one scale per 32 columns)
//   Q4_K   5120 B  nibbles    [2][128][16] + fp16 (d*sc, dmin*m) [128][2][2]
//   Q6_K   7168 B  low nibbles[2][128][16] + high 2-bit plane [128][16] + fp16 d*sc [128][4]
// The integer codes are the file's codes unchanged; the two-level K-quant scales are folded to one fp16
// factor per group at load.  Tiles are ordered [n_tile][k_block].
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

namespace ace {

enum WType : int { W_F16 = 0, W_Q8_0 = 1, W_Q4_K = 2, W_Q6_K = 3 };

// ggml type ids used in GGUF files (third_party/ggml/include/ggml.h enum ggml_type)
enum GgmlType : int {
    GGML_F32 = 0, GGML_F16 = 1, GGML_Q8_0 = 8, GGML_Q4_K = 12, GGML_Q6_K = 14, GGML_BF16 = 30,
    GGML_I8 = 24, GGML_I16 = 25, GGML_I32 = 26, GGML_I64 = 27, GGML_F64 = 28,
};

constexpr int TILE_N = 128;
constexpr int TILE_K = 64;

inline int wtype_tile_bytes(int wt) {
    switch (wt) {
        case W_F16: return 16384;
        case W_Q8_0: return 8704;
        case W_Q4_K: return 5120;
        case W_Q6_K: return 7168;
    }
    return 0;
}

// bytes of one row of n elements in a ggml type; 0 if n is not a whole number of blocks / unknown type
size_t ggml_row_bytes(int ggml_type, int64_t n);
int    ggml_block_elems(int ggml_type);
const char* ggml_type_name(int ggml_type);

// decode one row of `n` elements of a ggml type to fp32 (bit-exact with the ggml dequantisers)
bool dequant_row(int ggml_type, const void* src, float* dst, int64_t n);
// fp32 row -> Q8_0 blocks (ggml quantize_row_q8_0_ref: d = amax/127 as fp16, q = round(x/d))
void quantize_row_q8_0(const float* src, void* dst, int64_t n);
// fp32 row -> Q4_K / Q6_K super-blocks, bit-identical to ggml's quantize_row_q4_K_ref / quantize_row_q6_K_ref (n % 256 == 0)
void quantize_row_q4_k(const float* src, void* dst, int64_t n);
void quantize_row_q6_k(const float* src, void* dst, int64_t n);

float    f16_to_f32(uint16_t h);
uint16_t f32_to_f16(float f);  // round-to-nearest-even, same as the hardware cvt.rn.f16.f32
float    bf16_to_f32(uint16_t h);

struct PackedWeight {
    void*   dev = nullptr;   // device tiles
    void*   dev_f16 = nullptr; // quantised weights only: the same tiles decoded once to the F16 image (gemm_predecode);
                               // used when the GEMM has enough rows to be compute-bound (gemm.h: predecode_min_rows)
    int     wtype = W_F16;
    int     N = 0;           // output features (rows of W)
    int     K = 0;           // input features per tap
    int     taps = 1;
    int     kb_per_tap = 0;  // ceil(K/64)
    int     n_tiles = 0;     // ceil(N/128)
    int     tile_bytes = 0;
    size_t  bytes = 0;
    size_t  tile_offset(int nt, int kb) const {
        return (static_cast<size_t>(nt) * (static_cast<size_t>(taps) * kb_per_tap) + kb) * tile_bytes;
    }
};

// Build the host image of a packed weight.
//  * pack_tiles_f32: W given as fp32 [N, taps, K] (row-major) -> F16 tiles (values rounded to fp16), or,
//    with wtype == W_Q8_0 / W_Q4_K / W_Q6_K and taps == 1 and K a whole number of blocks, quantised on the fly like the
//    reference's ACE_GGML_*_WEIGHT_QTYPE path (acestep_dit_model.cpp:156-192).
//  * pack_tiles_ggml: W given as raw ggml rows [N][row_bytes] of type Q8_0 / Q4_K / Q6_K, K % 64 == 0 (K % 256
//    for the K-quants); the codes are carried over unchanged.
bool pack_tiles_f32(const float* w, int N, int taps, int K, int wtype, std::vector<uint8_t>& out, PackedWeight& meta,
                    std::string& err);
bool pack_tiles_ggml(const void* rows, int ggml_type, int N, int K, std::vector<uint8_t>& out, PackedWeight& meta,
                     std::string& err);

// Host emulation of the kernel's tile decoder (same bit manipulation, float arithmetic): decodes the packed
// image back to fp32 [N, taps*K]. Test aid: pack -> unpack must reproduce dequant_row up to the fp16 rounding
// of the folded scales.
void unpack_tiles_host(const std::vector<uint8_t>& img, const PackedWeight& meta, std::vector<float>& w);

}  // namespace ace
