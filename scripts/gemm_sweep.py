#!/usr/bin/env python
"""Tile-shape / split-K sweep of the tap-GEMM on the DiT and VAE problem shapes (tuning aid, not a bench line).
Times come from CUDA events around the single kernel launch inside ace_b200_debug_gemm."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ace-step-1.5-ggml_b200"), os.path.join(ROOT, "tests")]

from acestep_b200.bridge import B200Bridge  # noqa: E402
from blocks import random_rows  # noqa: E402


def main():
    br = B200Bridge()
    g = np.random.default_rng(0)
    shapes = [(750, 4096, 2048), (750, 2048, 2048), (750, 12288, 2048), (750, 2048, 6144),
              (3000, 4096, 2048), (3000, 2048, 2048), (3000, 12288, 2048), (3000, 2048, 6144),
              (6000, 4096, 2048), (6000, 2048, 2048), (6000, 12288, 2048), (6000, 2048, 6144)]
    wtypes = [0, 12] if len(sys.argv) < 2 else [int(x) for x in sys.argv[1].split(",")]
    for wt in wtypes:
        for (M, N, K) in shapes:
            A = g.standard_normal((1, M, K), dtype=np.float32)
            W = (g.standard_normal((N, K), dtype=np.float32) * 0.02) if wt == 0 else random_rows(wt, N, K, seed=1)
            R = np.zeros((1, M, N), dtype=np.float32)
            fl = 2.0 * M * N * K
            res = []
            for bn in (128, 256, 512):  # 512 = the CTA-pair kernel (256 x 256 per cluster)
                for sk in (1, 2, 3, 4):
                    os.environ["ACE_B200_GEMM_2CTA"] = "1" if bn == 512 else "0"
                    os.environ["ACE_B200_GEMM_BN"] = str(min(bn, 256))
                    os.environ["ACE_B200_GEMM_SPLITK"] = str(sk)
                    best = 1e9
                    for rep in range(3):
                        br.debug_gemm(A, W, w_type=wt, N=N, resid=R)
                        best = min(best, br.last_device_ms())
                    res.append(f"bn{bn}/sk{sk}: {best * 1e3:6.1f}us {fl / best / 1e9:6.0f}TF")
            print(f"wt={wt} M={M} N={N} K={K}: " + " | ".join(res), flush=True)
    br.close()


if __name__ == "__main__":
    main()
