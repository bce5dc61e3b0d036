#!/usr/bin/env python
"""Per-shape timing of the DiT's GEMMs on F16 tiles vs packed quantised tiles decoded in the kernel (default routes), at the row
counts of the BASELINE configs (tuning aid, not a bench line).  CUDA events around the single launch inside ace_b200_debug_gemm."""
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
    Ms = [int(x) for x in (sys.argv[1].split(",") if len(sys.argv) > 1 else ["750", "3000", "12000"])]
    wts = [int(x) for x in (sys.argv[2].split(",") if len(sys.argv) > 2 else ["0", "12", "8"])]
    shapes = [("qkv", 4096, 2048, {}), ("o_proj", 2048, 2048, {"resid": True}), ("gate_up", 12288, 2048, {"swiglu": True}), ("down", 2048, 6144, {"resid": True})]
    for M in Ms:
        for name, N, K, kw in shapes:
            A = g.standard_normal((1, M, K), dtype=np.float32)
            out = []
            for wt in wts:
                W = (g.standard_normal((N, K), dtype=np.float32) * 0.02) if wt == 0 else random_rows(wt, N, K, seed=1)
                args = dict(kw)
                if args.pop("resid", False):
                    args["resid"] = np.zeros((1, M, N), dtype=np.float32)
                best = 1e9
                for rep in range(4):
                    br.debug_gemm(A, W, w_type=wt, N=N, **args)
                    best = min(best, br.last_device_ms())
                out.append(f"wt{wt}: {best * 1e3:7.1f} us {2.0 * M * N * K / best / 1e9:6.0f} TF")
            print(f"M={M:6d} {name:8s} N={N:5d} K={K:4d} | " + " | ".join(out), flush=True)
    br.close()


if __name__ == "__main__":
    main()
