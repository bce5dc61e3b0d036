// Element-wise, normalisation and small reduction kernels of the DiT / VAE path (see kernels.h).
// All are HBM-bound: vectorised 16-byte accesses, warp-shuffle reductions, one pass over the data.
#include "kernels.h"

#include <cmath>

namespace ace {

namespace {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

inline unsigned blocks_for(long n, int threads) { return (unsigned)((n + threads - 1) / threads); }

// ---------------------------------------------------------------- pack
__global__ void pack_dit_input_kernel(const float* __restrict__ hidden, const float* __restrict__ context,
                                      __half* __restrict__ x0, int B, int T, int S, int patch, int ctx_dim,
                                      int audio_dim) {
    const int in_ch = ctx_dim + audio_dim;
    const long total = (long)B * S * patch * in_ch;
    const long idx = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const int c = (int)(idx % in_ch);
    long r = idx / in_ch;
    const int k = (int)(r % patch);
    r /= patch;
    const int p = (int)(r % S);
    const int b = (int)(r / S);
    const int t = p * patch + k;
    float v = 0.f;
    if (t < T) {
        if (c < ctx_dim) v = context ? context[((long)b * T + t) * ctx_dim + c] : 0.f;
        else v = hidden ? hidden[((long)b * T + t) * audio_dim + (c - ctx_dim)] : 0.f;
    }
    x0[idx] = __float2half_rn(v);
}

__global__ void f32_to_f16_kernel(const float* __restrict__ src, __half* __restrict__ dst, long n) {
    const long i4 = ((long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i4 + 3 < n) {
        const float4 v = *reinterpret_cast<const float4*>(src + i4);
        __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
        uint2 o;
        o.x = *reinterpret_cast<uint32_t*>(&a);
        o.y = *reinterpret_cast<uint32_t*>(&b);
        *reinterpret_cast<uint2*>(dst + i4) = o;
    } else {
        for (long i = i4; i < n; ++i) dst[i] = __float2half_rn(src[i]);
    }
}

__global__ void patch_mask_kernel(const int32_t* __restrict__ mask, int32_t* __restrict__ pm, int B, int T, int S,
                                  int patch) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= B * S) return;
    const int b = idx / S, p = idx % S;
    int v = 0;
    if (!mask) v = 1;
    else
        for (int k = 0; k < patch; ++k) {
            const int t = p * patch + k;
            if (t < T && mask[(long)b * T + t] != 0) v = 1;
        }
    pm[idx] = v;
}

__global__ void timestep_features_kernel(const float* __restrict__ t, __half* __restrict__ out, int n) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n * 128) return;
    const int r = idx / 128, i = idx % 128;
    const float ts = t[r] * 1000.0f;
    const float f = expf(-logf(10000.0f) * (float)i / 128.0f);
    const float arg = ts * f;
    out[(long)r * 256 + i] = __float2half_rn(cosf(arg));
    out[(long)r * 256 + 128 + i] = __float2half_rn(sinf(arg));
}

__global__ void build_modulation_kernel(const float* __restrict__ tables, const float* __restrict__ proj,
                                        float* __restrict__ mod, const float* __restrict__ out_table,
                                        const float* __restrict__ temb_a, const float* __restrict__ temb_b,
                                        float* __restrict__ out_mod, int B, int layers, int H) {
    const long n_mod = (long)B * layers * 6 * H;
    const long n_out = (long)B * 2 * H;
    const long idx = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx < n_mod) {
        const int c = (int)(idx % H);
        long r = idx / H;
        const int j = (int)(r % 6);
        r /= 6;
        const int layer = (int)(r % layers);
        const int b = (int)(r / layers);
        float v = tables[((long)layer * 6 + j) * H + c] + proj[((long)b * 6 + j) * H + c];
        if (j == 1 || j == 4) v += 1.0f;
        mod[idx] = v;
    } else if (idx < n_mod + n_out) {
        const long k = idx - n_mod;
        const int c = (int)(k % H);
        const int j = (int)((k / H) % 2);
        const int b = (int)(k / (2 * H));
        float v = out_table[(long)j * H + c] + temb_a[(long)b * H + c] + (temb_b ? temb_b[(long)b * H + c] : 0.f);
        if (j == 1) v += 1.0f;
        out_mod[k] = v;
    }
}

// ---------------------------------------------------------------- RMSNorm (+ AdaLN modulate): one 128-thread CTA per row
// (rows x 4 warps keeps every SM busy at S = 375..3000; the row stays in registers between the two passes)
__global__ void __launch_bounds__(128)
rmsnorm_mod_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ scale1,
                   const float* __restrict__ shift, long mod_bstride, __half* __restrict__ y, long rows, int rows_per_b,
                   int H, float eps) {
    __shared__ float red[4];
    const long row = blockIdx.x;
    const int tid = threadIdx.x;
    const float* xr = x + row * H;
    const int b = (int)(row / rows_per_b);
    constexpr int MAXV = 4;  // up to H = 2048 held in registers; wider rows re-read from L1/L2
    float4 v[MAXV];
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int c = (i * 128 + tid) * 4;
        if (c < H) {
            v[i] = *reinterpret_cast<const float4*>(xr + c);
            ss += v[i].x * v[i].x + v[i].y * v[i].y + v[i].z * v[i].z + v[i].w * v[i].w;
        }
    }
    for (int c = (MAXV * 128 + tid) * 4; c < H; c += 512) {
        const float4 t = *reinterpret_cast<const float4*>(xr + c);
        ss += t.x * t.x + t.y * t.y + t.z * t.z + t.w * t.w;
    }
    ss = warp_sum(ss);
    if ((tid & 31) == 0) red[tid >> 5] = ss;
    __syncthreads();
    ss = red[0] + red[1] + red[2] + red[3];
    const float inv = 1.0f / sqrtf(ss / (float)H + eps);
    const float* s1 = scale1 ? scale1 + (long)b * mod_bstride : nullptr;
    const float* sh = shift ? shift + (long)b * mod_bstride : nullptr;
    __half* yr = y + row * H;
    auto emit = [&](int c, float4 t) {
        const float4 ww = *reinterpret_cast<const float4*>(w + c);
        float o0 = t.x * inv * ww.x, o1 = t.y * inv * ww.y, o2 = t.z * inv * ww.z, o3 = t.w * inv * ww.w;
        if (s1) {
            const float4 a = *reinterpret_cast<const float4*>(s1 + c);
            o0 *= a.x; o1 *= a.y; o2 *= a.z; o3 *= a.w;
        }
        if (sh) {
            const float4 a = *reinterpret_cast<const float4*>(sh + c);
            o0 += a.x; o1 += a.y; o2 += a.z; o3 += a.w;
        }
        __half2 h0 = __floats2half2_rn(o0, o1), h1 = __floats2half2_rn(o2, o3);
        uint2 o;
        o.x = *reinterpret_cast<uint32_t*>(&h0);
        o.y = *reinterpret_cast<uint32_t*>(&h1);
        *reinterpret_cast<uint2*>(yr + c) = o;
    };
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
        const int c = (i * 128 + tid) * 4;
        if (c < H) emit(c, v[i]);
    }
    for (int c = (MAXV * 128 + tid) * 4; c < H; c += 512) emit(c, *reinterpret_cast<const float4*>(xr + c));
}

// ---------------------------------------------------------------- per-head RMSNorm + NEOX RoPE, in place (head_dim 128)
// one warp per (row, head); lane owns dims {2l, 2l+1} and their rotate-half partners {64+2l, 64+2l+1}
__global__ void headnorm_rope_kernel(__half* __restrict__ buf, long row_stride, int col0, int n_heads,
                                     const float* __restrict__ w, const float2* __restrict__ rope_tab, long rows,
                                     int rows_per_b, float eps) {
    const long wid = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (wid >= rows * n_heads) return;
    const int lane = threadIdx.x & 31;
    const long row = wid / n_heads;
    const int head = (int)(wid % n_heads);
    __half* p = buf + row * row_stride + col0 + head * 128;
    const __half2 a2 = *reinterpret_cast<const __half2*>(p + 2 * lane);
    const __half2 b2 = *reinterpret_cast<const __half2*>(p + 64 + 2 * lane);
    float a0 = __low2float(a2), a1 = __high2float(a2), b0 = __low2float(b2), b1 = __high2float(b2);
    float ss = warp_sum(a0 * a0 + a1 * a1 + b0 * b0 + b1 * b1);
    const float inv = 1.0f / sqrtf(ss / 128.0f + eps);
    a0 *= inv * w[2 * lane];
    a1 *= inv * w[2 * lane + 1];
    b0 *= inv * w[64 + 2 * lane];
    b1 *= inv * w[64 + 2 * lane + 1];
    if (rope_tab) {
        const int pos = (int)(row % rows_per_b);
        const float2 cs0 = rope_tab[(long)pos * 64 + 2 * lane];
        const float2 cs1 = rope_tab[(long)pos * 64 + 2 * lane + 1];
        const float r0 = a0 * cs0.x - b0 * cs0.y, s0 = a0 * cs0.y + b0 * cs0.x;
        const float r1 = a1 * cs1.x - b1 * cs1.y, s1 = a1 * cs1.y + b1 * cs1.x;
        a0 = r0; b0 = s0; a1 = r1; b1 = s1;
    }
    *reinterpret_cast<__half2*>(p + 2 * lane) = __floats2half2_rn(a0, a1);
    *reinterpret_cast<__half2*>(p + 64 + 2 * lane) = __floats2half2_rn(b0, b1);
}

__global__ void rope_table_kernel(float2* __restrict__ tab, int S, int half_dim, float theta) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= S * half_dim) return;
    const int pos = idx / half_dim, i = idx % half_dim;
    // angle = pos * theta^(-2i/n_dims), evaluated in double so the table is exact to fp32 rounding
    const double ang = (double)pos * pow((double)theta, -(double)i / (double)half_dim);
    tab[idx] = make_float2((float)cos(ang), (float)sin(ang));
}

__global__ void euler_step_kernel(float* __restrict__ x, const float* __restrict__ vc, const float* __restrict__ vu,
                                  float g, float dt, long n) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float v = vc[i];
    if (vu) v = vu[i] + g * (v - vu[i]);
    x[i] -= v * dt;
}

// ---------------------------------------------------------------- tiny-C_out conv (VAE output conv)
// one warp per output position: lanes split the input channels, 16-byte loads, shuffle reduction
template <int COUT>
__global__ void conv_small_out_kernel(const __half* __restrict__ xs, const __half* __restrict__ w,
                                      const float* __restrict__ bias, float* __restrict__ out, int B, long L, int C,
                                      int taps, int pad) {
    const long wid = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (wid >= (long)B * L) return;
    const int lane = threadIdx.x & 31;
    const int b = (int)(wid / L);
    const long l = wid % L;
    float acc[COUT];
#pragma unroll
    for (int o = 0; o < COUT; ++o) acc[o] = 0.f;
    for (int tap = 0; tap < taps; ++tap) {
        const long ls = l + tap - pad;
        if (ls < 0 || ls >= L) continue;
        const __half* xr = xs + ((long)b * L + ls) * C;
        for (int c = lane * 8; c < C; c += 256) {
            const uint4 xv = *reinterpret_cast<const uint4*>(xr + c);
            const __half2* xh = reinterpret_cast<const __half2*>(&xv);
#pragma unroll
            for (int o = 0; o < COUT; ++o) {
                const uint4 wv = *reinterpret_cast<const uint4*>(w + ((long)o * taps + tap) * C + c);
                const __half2* wh = reinterpret_cast<const __half2*>(&wv);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float2 xf = __half22float2(xh[i]), wf = __half22float2(wh[i]);
                    acc[o] = fmaf(xf.x, wf.x, acc[o]);
                    acc[o] = fmaf(xf.y, wf.y, acc[o]);
                }
            }
        }
    }
#pragma unroll
    for (int o = 0; o < COUT; ++o) acc[o] = warp_sum(acc[o]);
    if (lane == 0) {
#pragma unroll
        for (int o = 0; o < COUT; ++o) out[((long)b * L + l) * COUT + o] = acc[o] + (bias ? bias[o] : 0.f);
    }
}

}  // namespace

void launch_pack_dit_input(const float* hidden, const float* context, __half* x0, int B, int T, int S, int patch,
                           int ctx_dim, int audio_dim, cudaStream_t st) {
    note_launch();
    ProfScope prof_scope(PROF_ELEM, 0.0, 0.0, st);
    const long total = (long)B * S * patch * (ctx_dim + audio_dim);
    pack_dit_input_kernel<<<blocks_for(total, 256), 256, 0, st>>>(hidden, context, x0, B, T, S, patch, ctx_dim, audio_dim);
}
void launch_f32_to_f16(const float* src, __half* dst, long n, cudaStream_t st) {
    note_launch();
    ProfScope prof_scope(PROF_ELEM, 0.0, 0.0, st);
    if (n <= 0) return;
    f32_to_f16_kernel<<<blocks_for((n + 3) / 4, 256), 256, 0, st>>>(src, dst, n);
}
void launch_patch_mask(const int32_t* mask, int32_t* pm, int B, int T, int S, int patch, cudaStream_t st) {
    note_launch();
    ProfScope prof_scope(PROF_ELEM, 0.0, 0.0, st);
    patch_mask_kernel<<<blocks_for((long)B * S, 256), 256, 0, st>>>(mask, pm, B, T, S, patch);
}
void launch_timestep_features(const float* t, __half* out, int n, cudaStream_t st) {
    note_launch();
    ProfScope prof_scope(PROF_ELEM, 0.0, 0.0, st);
    timestep_features_kernel<<<blocks_for((long)n * 128, 128), 128, 0, st>>>(t, out, n);
}
void launch_build_modulation(const float* tables, const float* proj, float* mod, const float* out_table,
                             const float* temb_a, const float* temb_b, float* out_mod, int B, int layers, int H,
                             cudaStream_t st) {
    note_launch();
    ProfScope prof_scope(PROF_ELEM, 0.0, 0.0, st);
    const long total = (long)B * layers * 6 * H + (long)B * 2 * H;
    build_modulation_kernel<<<blocks_for(total, 256), 256, 0, st>>>(tables, proj, mod, out_table, temb_a, temb_b, out_mod,
                                                                    B, layers, H);
}
void launch_rmsnorm_mod(const float* x, const float* w, const float* scale1, const float* shift, long mod_bstride,
                        __half* y, int B, int rows_per_b, int H, float eps, cudaStream_t st) {
    note_launch();
    ProfScope prof_scope(PROF_ELEM, 0.0, 0.0, st);
    const long rows = (long)B * rows_per_b;
    rmsnorm_mod_kernel<<<(unsigned)rows, 128, 0, st>>>(x, w, scale1, shift, mod_bstride, y, rows, rows_per_b, H, eps);
}
void launch_headnorm_rope(__half* buf, long row_stride, int col0, int n_heads, const float* w, const float2* rope_tab,
                          int B, int rows_per_b, float eps, cudaStream_t st) {
    note_launch();
    ProfScope prof_scope(PROF_ELEM, 0.0, 0.0, st);
    const long rows = (long)B * rows_per_b;
    headnorm_rope_kernel<<<blocks_for(rows * n_heads, 8), 256, 0, st>>>(buf, row_stride, col0, n_heads, w, rope_tab, rows,
                                                                        rows_per_b, eps);
}
void launch_build_rope_table(float2* tab, int S, int half_dim, float theta, cudaStream_t st) {
    note_launch();
    ProfScope prof_scope(PROF_ELEM, 0.0, 0.0, st);
    rope_table_kernel<<<blocks_for((long)S * half_dim, 256), 256, 0, st>>>(tab, S, half_dim, theta);
}
void launch_euler_step(float* x, const float* vc, const float* vu, float g, float dt, long n, cudaStream_t st) {
    note_launch();
    ProfScope prof_scope(PROF_ELEM, 0.0, 0.0, st);
    euler_step_kernel<<<blocks_for(n, 256), 256, 0, st>>>(x, vc, vu, g, dt, n);
}
void launch_conv_small_out(const __half* xs, const __half* w, const float* bias, float* out, int B, long L, int C,
                           int Cout, int taps, int pad, cudaStream_t st) {
    note_launch();
    ProfScope prof_scope(PROF_VAE_TAIL, 2.0 * B * L * C * Cout * taps, (double)B * L * (C * 2.0 + Cout * 4.0), st);
    const long warps = (long)B * L;
    if (Cout == 1) conv_small_out_kernel<1><<<blocks_for(warps, 8), 256, 0, st>>>(xs, w, bias, out, B, L, C, taps, pad);
    else if (Cout == 2) conv_small_out_kernel<2><<<blocks_for(warps, 8), 256, 0, st>>>(xs, w, bias, out, B, L, C, taps, pad);
}

}  // namespace ace
