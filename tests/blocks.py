"""Random ggml quantised rows for kernel-level tests: any bit pattern is a valid block, so random codes with
sane fp16 scales exercise every field of the decoders (layouts: ggml-common.h:219-223, 291-305, 326-335)."""
import numpy as np


def _f16_bytes(x):
    return np.asarray(x, dtype=np.float16).view(np.uint8)


def random_rows(ggml_type: int, N: int, K: int, seed: int = 0, scale: float = 0.02) -> np.ndarray:
    g = np.random.default_rng(seed)
    if ggml_type == 8:  # q8_0: half d; int8 qs[32]
        nb = K // 32
        out = np.empty((N, nb, 34), dtype=np.uint8)
        d = (g.random((N, nb)) * scale / 64 + 1e-4).astype(np.float16)
        out[:, :, 0:2] = d.view(np.uint8).reshape(N, nb, 2)
        out[:, :, 2:] = g.integers(0, 256, (N, nb, 32), dtype=np.uint8)
        return out.reshape(N, nb * 34)
    if ggml_type == 12:  # q4_K: half d, dmin; u8 scales[12]; u8 qs[128]
        nb = K // 256
        out = np.empty((N, nb, 144), dtype=np.uint8)
        d = (g.random((N, nb)) * scale / 256 + 1e-5).astype(np.float16)
        dm = (g.random((N, nb)) * scale / 64 + 1e-5).astype(np.float16)
        out[:, :, 0:2] = d.view(np.uint8).reshape(N, nb, 2)
        out[:, :, 2:4] = dm.view(np.uint8).reshape(N, nb, 2)
        out[:, :, 4:] = g.integers(0, 256, (N, nb, 140), dtype=np.uint8)
        return out.reshape(N, nb * 144)
    if ggml_type == 14:  # q6_K: u8 ql[128]; u8 qh[64]; i8 scales[16]; half d
        nb = K // 256
        out = np.empty((N, nb, 210), dtype=np.uint8)
        out[:, :, :208] = g.integers(0, 256, (N, nb, 208), dtype=np.uint8)
        d = (g.random((N, nb)) * scale / 2048 + 1e-6).astype(np.float16)
        out[:, :, 208:210] = d.view(np.uint8).reshape(N, nb, 2)
        return out.reshape(N, nb * 210)
    raise ValueError(ggml_type)
