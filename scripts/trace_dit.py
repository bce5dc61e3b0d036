#!/usr/bin/env python
"""Phase traces of every tap-GEMM of one DiT forward on the bench workload (ACE_B200_GEMM_TRACE=1): prints the launches of layer 1."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ace-step-1.5-ggml_b200"), os.path.join(ROOT, "tests")]
import bench  # noqa: E402
from parity_common import full_vae_dir, turbo_dit_dir  # noqa: E402
from acestep_b200.bridge import B200Bridge  # noqa: E402

T = int(sys.argv[1]) if len(sys.argv) > 1 else 1500
br = B200Bridge()
br.load_dit(turbo_dit_dir("F16"))
noise, ctx, enc = bench.synth_inputs(T, 512)
for _ in range(3):
    br.dit_forward_tfirst(noise[0], ctx[0], enc[0], None, None, 1.0, 1.0)
print(f"forward T={T}: device {br.last_device_ms():.3f} ms", flush=True)
os.environ["ACE_B200_GEMM_TRACE"] = "1"
br.dit_forward_tfirst(noise[0], ctx[0], enc[0], None, None, 1.0, 1.0)
br.close()
