// Oobleck VAE decoder: weight-norm folded conv weights as packed F16 tiles + the fused decode schedule.
// reference: acestep_vae_model.h:13-78 (Config / handles), acestep_vae_model.cpp:56-126 (config),
//            503-641 (snake / conv / residual-unit / decoder-block loaders), 761-955 (names), 957-1002 (forward_decode)
#pragma once
#include <string>
#include <vector>

#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include "wpack.h"

namespace ace {

struct VaeConfig {
    int audio_channels = 0, encoder_hidden_size = 0, decoder_channels = 0, decoder_input_channels = 0;
    int sampling_rate = 0;
    std::vector<int> downsampling_ratios, channel_multiples, upsampling_ratios;
    int hop_length = 1;
};

struct VaeSnake {
    float* alpha = nullptr;     // exp(alpha) [C]
    float* inv_beta = nullptr;  // 1 / exp(beta) [C]
};

struct VaeConv {
    PackedWeight w;
    float* bias = nullptr;  // [N of the GEMM] (conv-transpose: tiled over the stride phases)
    int c_in = 0, c_out = 0, k = 1, dilation = 1, stride = 1, pad = 0;
};

struct VaeResUnit {
    VaeSnake snake1, snake2;
    VaeConv conv1, conv2;
};

struct VaeBlock {
    VaeSnake snake1;
    VaeConv conv_t;
    VaeResUnit res[3];
};

// encoder block: 3 residual units (dilations 1, 3, 9) at C_in, Snake, strided conv k = 2*stride (acestep_vae_model.cpp:643-668)
struct VaeEncBlock {
    VaeResUnit res[3];
    VaeSnake snake1;
    VaeConv conv;  // packed as a 2-tap GEMM over rows of `stride` consecutive positions: K = stride * C_in per tap
};

struct VaeModel {
    VaeConfig cfg;
    VaeConv conv1;
    std::vector<VaeBlock> blocks;
    // encoder (acestep_vae_model.cpp:913-929, 1004-1044); present when the file holds the encoder.* tensors
    bool has_encoder = false;
    VaeConv enc_conv1;             // k7, audio channels zero-padded to 8 input channels
    std::vector<VaeEncBlock> enc_blocks;
    VaeSnake enc_snake;
    VaeConv enc_conv2;             // k3 -> encoder_hidden_size (mean | scale)
    VaeSnake snake_out;
    __half* conv2_w = nullptr;  // [audio_channels, 7, C] fp16 (tiny-N tail conv)
    int conv2_cin = 0;
    std::vector<void*> allocs;
    size_t weight_bytes = 0;
    void release();
};

int vae_load(const std::string& model_dir, VaeModel& m, std::string& err);

struct VaeWorkspace {
    size_t cap_frames = 0;
    int capB = 0;
    float* d_lat = nullptr;   // [B, T, 64] f32
    __half* lat16 = nullptr;
    float* x = nullptr;       // fp32 residual stream (largest level)
    __half* xs = nullptr;     // snake'd fp16 copy feeding the next conv
    __half* h = nullptr;      // fp16 intermediate inside a residual unit / conv-transpose input
    std::vector<void*> allocs;
    void release();
};

bool vae_workspace_reserve(const VaeModel& m, VaeWorkspace& ws, int B, int n_frames, std::string& err);

// number of latent frames the encoder produces for n_samples audio samples (floor at every strided conv, PyTorch rule)
long vae_encode_out_frames(const VaeModel& m, long n_samples);
// audio (device, [n_samples, audio_channels] fp32 interleaved) -> posterior mean latents (device, [frames, latent_ch] fp32).
// One clip per call (the reference ABI's shape, acestep_ggml.cpp:975-1072).  Scratch comes from `ws` (grown on demand).
struct VaeEncWorkspace {
    size_t cap_samples = 0;
    __half* a16 = nullptr;   // padded audio
    float* x = nullptr;      // fp32 residual stream
    __half *xs = nullptr, *h = nullptr, *xp = nullptr;
    float* y = nullptr;      // conv2 output [frames, encoder_hidden_size]
    std::vector<void*> allocs;
    void release();
};
bool vae_encode_workspace_reserve(const VaeModel& m, VaeEncWorkspace& ws, long n_samples, std::string& err);
cudaError_t vae_encode_device(const VaeModel& m, VaeEncWorkspace& ws, const float* d_audio, long n_samples, float* d_latents,
                              cudaStream_t st);

// latents in ws.d_lat [B, n_frames, latent_ch] -> audio [B, n_frames*hop, audio_ch] (device pointer `audio`)
// `win` (optional, B = 1): keep only the samples of latent frames [frame_lo, frame_hi) of this window and store them at sample
// offset dst_frame * hop of `audio` (the assembled clip, possibly peer memory) instead of in place.
struct VaeOutWindow {
    long frame_lo = 0, frame_hi = 0, dst_frame = 0;
};
cudaError_t vae_decode_device(const VaeModel& m, VaeWorkspace& ws, int B, int n_frames, float* audio, cudaStream_t st,
                              const VaeOutWindow* win = nullptr);

// One residual unit at 128 channels in ONE launch (vae_resunit.cu): k7 dilated conv -> Snake -> k1 conv -> + x -> next Snake, the
// intermediate h staying in shared memory.  xs_in: Snake'd input [B, L, 128] fp16; x_in: residual stream (fp16); x_out (may be
// null, may alias x_in): new residual stream; xs_out (must not alias xs_in: neighbouring tiles read its halo): next Snake'd copy.
bool vae_resunit128_supported(const VaeResUnit& ru);
cudaError_t launch_vae_resunit128(const __half* xs_in, const __half* x_in, __half* x_out, __half* xs_out, int B, long L,
                                  const VaeResUnit& ru, const VaeSnake& next, cudaStream_t st);

}  // namespace ace
