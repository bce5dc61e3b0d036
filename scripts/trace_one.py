#!/usr/bin/env python
"""Phase trace of single GEMM launches (ACE_B200_GEMM_TRACE=1: per-CTA clock64 stamps, printed by the library)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ace-step-1.5-ggml_b200"), os.path.join(ROOT, "tests")]
from acestep_b200.bridge import B200Bridge  # noqa: E402

br = B200Bridge()
g = np.random.default_rng(0)
os.environ["ACE_B200_GEMM_2CTA"] = "0"
for (M, N, K, bn, sk) in [(750, 2048, 2048, 128, 1), (750, 2048, 2048, 128, 3), (750, 2048, 2048, 256, 2), (750, 4096, 2048, 256, 1),
                          (750, 12288, 2048, 256, 1), (750, 2048, 6144, 256, 3), (6000, 2048, 2048, 256, 1)]:
    A = g.standard_normal((1, M, K), dtype=np.float32)
    W = g.standard_normal((N, K), dtype=np.float32) * 0.02
    R = np.zeros((1, M, N), dtype=np.float32)
    os.environ["ACE_B200_GEMM_BN"] = str(bn)
    os.environ["ACE_B200_GEMM_SPLITK"] = str(sk)
    os.environ.pop("ACE_B200_GEMM_TRACE", None)
    best = 1e9
    for _ in range(4):
        br.debug_gemm(A, W, w_type=0, N=N, resid=R)
        best = min(best, br.last_device_ms())
    print(f"M={M} N={N} K={K} bn={bn} sk={sk}: {best * 1e3:.1f} us", flush=True)
    os.environ["ACE_B200_GEMM_TRACE"] = "1"
    br.debug_gemm(A, W, w_type=0, N=N, resid=R)
    if sk == 1:  # same launch with a plain read-add-write epilogue instead of red.add
        os.environ["ACE_B200_GEMM_ATOMIC"] = "0"
        os.environ.pop("ACE_B200_GEMM_TRACE", None)
        best = 1e9
        for _ in range(4):
            br.debug_gemm(A, W, w_type=0, N=N, resid=R)
            best = min(best, br.last_device_ms())
        print(f"  non-atomic: {best * 1e3:.1f} us", flush=True)
        os.environ["ACE_B200_GEMM_TRACE"] = "1"
        br.debug_gemm(A, W, w_type=0, N=N, resid=R)
        os.environ.pop("ACE_B200_GEMM_ATOMIC", None)
# SwiGLU epilogue (gate_up projection): no residual, fp16 output of N/2 columns
os.environ.pop("ACE_B200_GEMM_SPLITK", None)
os.environ["ACE_B200_GEMM_BN"] = "256"
A = g.standard_normal((1, 750, 2048), dtype=np.float32)
W = g.standard_normal((12288, 2048), dtype=np.float32) * 0.02
os.environ.pop("ACE_B200_GEMM_TRACE", None)
best = 1e9
for _ in range(4):
    br.debug_gemm(A, W, w_type=0, N=12288, swiglu=True)
    best = min(best, br.last_device_ms())
print(f"SwiGLU M=750 N=12288 K=2048: {best * 1e3:.1f} us", flush=True)
os.environ["ACE_B200_GEMM_TRACE"] = "1"
br.debug_gemm(A, W, w_type=0, N=12288, swiglu=True)
br.close()
