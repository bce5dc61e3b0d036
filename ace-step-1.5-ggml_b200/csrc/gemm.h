// Tap-GEMM on sm_100a: the one dense-contraction kernel of the path.
//
//   D[b, l, n] = sum_{tap, c}  A[b, l + tap*tap_stride + tap_offset, c] * W[n, tap, c]
//
// A is an fp16 activation tensor [B, L, C] (rows outside [0,L) read as zero: TMA out-of-bounds fill),
// W is a packed weight (wpack.h) in F16 / Q8_0 / Q4_K / Q6_K codes that the kernel dequantises to fp16
// in shared memory, the accumulator is fp32 in TMEM.  One kernel covers
//   * every ggml_mul_mat of the DiT graph (taps = 1)                 reference: acestep_dit_model.cpp:1194-1196,1257,1295-1302,1381,1412,1528-1531,1551
//   * ggml_conv_1d of the VAE (taps = 7 or 1, stride = dilation)      reference: acestep_vae_model.cpp:709
//   * ggml_conv_transpose_1d (taps = 2, N = stride*C_out phases)      reference: acestep_vae_model.cpp:696-707
// with the element-wise neighbours (bias, AdaLN gate, residual add, SwiGLU, Snake) fused in the epilogue.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include "wpack.h"

namespace ace {

enum EpiMode : int {
    EPI_STORE  = 0,  // v = acc (+bias) (*gate) (+resid) -> out_f32 / out_f16 (optionally snake'd)
    EPI_SWIGLU = 1,  // packed rows interleave (gate_j, up_j): out_f16[m, j] = silu(acc[2j]) * acc[2j+1]
    // Q/K/V projections: columns are heads of 128; the first hn_na columns get per-head RMSNorm with weight hn_wa,
    // the next hn_nb columns with hn_wb, the rest are stored as they are; normed heads are then rotated (NEOX RoPE,
    // pairs (i, i+64), angle table `rope`[l]) when rope != null.  out_f16 only.
    // reference: acestep_dit_model.cpp:1202-1210 (rms_norm(q/k) + ggml_rope_ext)
    EPI_HEADNORM = 2,
};

struct GemmEpilogue {
    int          mode          = EPI_STORE;
    const float* bias          = nullptr;  // [N]
    const float* gate          = nullptr;  // [B, gate_bstride] per-sample column scale (AdaLN gate)
    long         gate_bstride  = 0;
    const float* resid         = nullptr;  // fp32, same indexing as out_f32 (may alias it)
    const __half* resid16      = nullptr;  // 16-bit residual stream (VAE): same indexing; exclusive with resid
    float*       out_f32       = nullptr;
    __half*      out_x16       = nullptr;  // v (after bias / residual, BEFORE Snake) rounded to fp16: the 16-bit residual stream
    __half*      out_f16       = nullptr;
    long         ld_out        = 0;        // elements between consecutive l rows of the outputs
    long         out_bstride   = 0;        // elements between samples
    long         out_shift     = 0;        // flat element shift inside a sample (conv-transpose crop: -pad*C_out)
    long         out_valid     = 0;        // valid flat elements per sample (bounds check after the shift)
    const float* snake_alpha   = nullptr;  // [N] exp(alpha)      out_f16 = v + sin^2(a v) * inv_b
    const float* snake_invbeta = nullptr;  // [N] 1/exp(beta)
    int          snake_period  = 0;        // channel index = n % snake_period (0 = n)
    int          atomic        = 0;        // out_f32 += v via red.add (split-K); bias only from split 0
    int          silu_f16      = 0;        // out_f16 = silu(v) (time-embedding MLP), out_f32 keeps v
    const float* hn_wa         = nullptr;  // EPI_HEADNORM: [128] norm weight of the first hn_na columns
    const float* hn_wb         = nullptr;  //               [128] norm weight of the next hn_nb columns
    int          hn_na = 0, hn_nb = 0;
    const float2* rope         = nullptr;  // [rows_per_sample, 64] (cos, sin)
    float        hn_eps        = 1e-6f;
};

struct GemmProblem {
    const __half* A = nullptr;
    int  B = 1, L = 0, C = 0;      // activation tensor [B, L, C]
    long a_lstride = 0;            // elements between rows (>= C, multiple of 8)
    long a_bstride = 0;            // elements between samples
    int  l_out = 0;                // rows of D per sample (L for linear/conv, L+1 for conv-transpose)
    int  tap_stride = 0, tap_offset = 0;
    const PackedWeight* W = nullptr;
    int  split_k = 1;              // 1 = never split K, 0 = automatic (only in-place accumulating epilogues), n = force
    GemmEpilogue epi;
};

// Launches on `stream`; returns cudaSuccess or the failing error. `simt` selects the plain-FMA debug kernel
// (same operands, same epilogue) that the tensor-core kernel is unit-tested against.
cudaError_t gemm_launch(const GemmProblem& p, cudaStream_t stream, bool simt = false);

// Quantised weights stay packed in HBM.  Three routes feed the tensor cores bit-identical fp16 operands (tests/test_kernels_gpu.py,
// test_golden_gpu.py, test_y_full_size_properties_gpu.py):
//   * launches of < decode_pass_min_rows() rows (default 2048): decoded IN THE GEMM — packed codes land in shared memory by TMA bulk copy,
//     decoder warps expand them to the fp16 UMMA tile, tcgen05 consumes it (single-CTA kernel for one row tile, CTA-pair kernel beyond;
//     the pair's decoders need a cluster-scoped proxy fence because the peer's tensor core reads their tile: gemm_sm100.cu);
//   * launches of >= decode_pass_min_rows() rows: DECODE PASS — the matrix of this launch is decoded by predecode_kernel into a <= 50 MB
//     scratch of the calling stream, stream-ordered in front of the GEMM, which runs the F16 CTA-pair kernel (ACE_B200_DECODE_PASS=0
//     forbids it, ACE_B200_DECODE_PASS_MIN_M moves the threshold);
//   * opt-in A/B route of round 1: ACE_B200_PREDECODE=1 decodes every quantised weight once at load (a resident F16 twin, +3.15 GB for the
//     turbo DiT) and GEMMs with >= predecode_min_rows() rows (ACE_B200_PREDECODE_MIN_M, default 256) read it through the F16 kernels.
size_t      gemm_predecode_bytes(const PackedWeight& w);
cudaError_t gemm_predecode(const PackedWeight& w, void* dst, cudaStream_t stream);
bool        predecode_enabled();
int         predecode_min_rows();
int         decode_pass_min_rows();   // 0 = decode pass forbidden
void        decode_scratch_release(cudaStream_t stream);   // frees the decode-pass scratch of a stream about to be destroyed

// true when ACE_B200_GEMM=simt is set (debug bisecting aid; never the default)
bool gemm_force_simt();

int  device_sm_count();

}  // namespace ace
