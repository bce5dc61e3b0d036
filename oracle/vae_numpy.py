"""TEST INFRASTRUCTURE (oracle). NumPy restatement of the reference's Oobleck VAE decode.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module.

What it restates (file:line in /root/reference/acestep_ggml/cpp unless noted):
  load_conv_weight_norm   acestep_vae_model.cpp:520-588  (w = v g / sqrt(sum_row v^2 + 1e-12), rows over d0; stored F16)
  snake_forward           acestep_vae_model.cpp:682-692  (x + sin^2(exp(alpha) x) / exp(beta))
  conv_forward            acestep_vae_model.cpp:694-712  (ggml_conv_1d = F16 im2col x F16 kernel, F32 accumulate,
                                                          ggml.c:4337-4426; conv_transpose_1d p0=0 then centre crop)
  residual_forward        acestep_vae_model.cpp:724-733
  decoder_block_forward   acestep_vae_model.cpp:735-742
  forward_decode          acestep_vae_model.cpp:957-1002
  tiling policy           acestep_ggml.cpp:2114-2229 (window = core +- overlap, trim, concatenate)

`f16=True` (default) rounds conv inputs and folded weights to fp16 exactly where the reference does; accumulation is
float64.  Pinned against golden vectors from the unmodified reference (tests/golden/make_golden.py).
"""
from __future__ import annotations

from typing import Dict, List

import numpy as np

F = np.float64


def _h(x, f16):
    return x.astype(np.float32).astype(np.float16).astype(F) if f16 else x.astype(F)


def fold_weight_norm(T: Dict[str, np.ndarray], prefix: str) -> np.ndarray:
    v = T[prefix + ".weight_v"].astype(np.float32)
    g = T[prefix + ".weight_g"].astype(np.float32).reshape(-1)
    ss = (v.astype(F) ** 2).reshape(v.shape[0], -1).sum(axis=1)
    scale = g / np.sqrt(ss.astype(np.float32) + np.float32(1e-12))
    return (v * scale[:, None, None]).astype(np.float32)


def snake(T, prefix, x):
    """x: [L, C]"""
    a = np.exp(T[prefix + ".alpha"].astype(np.float32).reshape(-1)).astype(F)
    b = np.exp(T[prefix + ".beta"].astype(np.float32).reshape(-1)).astype(F)
    return x + np.sin(a * x) ** 2 / b


def conv1d(x, w, bias, dilation: int, f16: bool):
    """x [L, Cin]; w [Cout, Cin, K] (PyTorch); 'same' padding (K-1)*dilation/2; returns [L, Cout]"""
    L = x.shape[0]
    Cout, _, K = w.shape
    pad = (K - 1) * dilation // 2
    xh = _h(x, f16)
    wh = _h(w, f16)
    xp = np.zeros((L + 2 * pad, x.shape[1]), dtype=F)
    xp[pad:pad + L] = xh
    y = np.zeros((L, Cout), dtype=F)
    for k in range(K):
        y += xp[k * dilation:k * dilation + L] @ wh[:, :, k].T
    return y if bias is None else y + bias.astype(F)


def conv_transpose1d(x, w, bias, stride: int, f16: bool):
    """x [L, Cin]; w [Cin, Cout, K] with K = 2*stride; full output then centre crop of ceil(stride/2) per side"""
    L = x.shape[0]
    _, Cout, K = w.shape
    pad = (stride + 1) // 2
    xh = _h(x, f16)
    wh = _h(w, f16)
    full = np.zeros(((L - 1) * stride + K, Cout), dtype=F)
    for k in range(K):
        full[k:k + (L - 1) * stride + 1:stride] += xh @ wh[:, :, k]
    target = (L - 1) * stride - 2 * pad + (K - 1) + 1
    start = (full.shape[0] - target) // 2
    y = full[start:start + target]
    return y if bias is None else y + bias.astype(F)


def decode(T: Dict[str, np.ndarray], cfg: dict, latents: np.ndarray, f16: bool = True, trace: List[np.ndarray] = None) -> np.ndarray:
    """latents [n_frames, latent_ch] -> audio [n_frames*hop, audio_ch] (float64)"""
    ups = list(reversed(cfg["downsampling_ratios"]))
    x = conv1d(latents.astype(F), fold_weight_norm(T, "decoder.conv1"), T["decoder.conv1.bias"], 1, f16)
    for i, s in enumerate(ups):
        p = f"decoder.block.{i}"
        x = snake(T, p + ".snake1", x)
        x = conv_transpose1d(x, fold_weight_norm(T, p + ".conv_t1"), T[p + ".conv_t1.bias"], s, f16)
        for j, dil in enumerate((1, 3, 9)):
            r = f"{p}.res_unit{j + 1}"
            y = conv1d(snake(T, r + ".snake1", x), fold_weight_norm(T, r + ".conv1"), T[r + ".conv1.bias"], dil, f16)
            y = conv1d(snake(T, r + ".snake2", y), fold_weight_norm(T, r + ".conv2"), T[r + ".conv2.bias"], 1, f16)
            x = x + y
        if trace is not None:
            trace.append(x.copy())
    x = snake(T, "decoder.snake1", x)
    return conv1d(x, fold_weight_norm(T, "decoder.conv2"), None, 1, f16)


def decode_tiled(T, cfg, latents, chunk: int = 128, overlap: int = 32, f16: bool = True) -> np.ndarray:
    """the reference's C++ tiling policy (acestep_ggml.cpp:2114-2229)"""
    hop = int(np.prod(cfg["downsampling_ratios"]))
    n = latents.shape[0]
    stride = max(1, chunk - 2 * overlap)
    parts = []
    for c0 in range(0, n, stride):
        c1 = min(c0 + stride, n)
        w0, w1 = max(0, c0 - overlap), min(n, c1 + overlap)
        a = decode(T, cfg, latents[w0:w1], f16)
        s0 = (c0 - w0) * hop
        parts.append(a[s0:s0 + (c1 - c0) * hop])
    return np.concatenate(parts, axis=0)
