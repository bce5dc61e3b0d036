#!/usr/bin/env python
"""Phase trace of the VAE's last-level convolutions as tap-GEMMs (ACE_B200_GEMM_TRACE=1)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ace-step-1.5-ggml_b200"), os.path.join(ROOT, "tests")]
from acestep_b200.bridge import B200Bridge  # noqa: E402

br = B200Bridge()
g = np.random.default_rng(0)
L = 148 * 128 * 16
for (C, N, taps, dil) in [(128, 128, 7, 1), (128, 128, 1, 1), (256, 256, 7, 1), (256, 256, 2, 1)]:
    A = g.standard_normal((1, L, C), dtype=np.float32)
    W = (g.standard_normal((N, taps, C), dtype=np.float32) * (taps * C) ** -0.5).astype(np.float32)
    bias = g.standard_normal((N,), dtype=np.float32)
    kw = dict(w_type=0, N=N, bias=bias, taps=taps, tap_stride=dil if taps > 1 else 0, tap_offset=-(taps // 2) * dil)
    os.environ.pop("ACE_B200_GEMM_TRACE", None)
    best = 1e9
    for _ in range(3):
        br.debug_gemm(A, W if taps > 1 else W[:, 0], **kw)
        best = min(best, br.last_device_ms())
    fl = 2.0 * L * N * C * taps
    print(f"conv C={C} N={N} taps={taps} dil={dil} L={L}: {best * 1e3:.1f} us  {fl / best / 1e9:.0f} TFLOP/s  out {L * N * 4 / best / 1e6:.0f} GB/s(f32 store)", flush=True)
    os.environ["ACE_B200_GEMM_TRACE"] = "1"
    os.environ["ACE_B200_GEMM_2CTA"] = "0"
    br.debug_gemm(A, W if taps > 1 else W[:, 0], **kw)
    os.environ.pop("ACE_B200_GEMM_2CTA", None)
br.close()
