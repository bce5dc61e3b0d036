#!/usr/bin/env python
"""Per-launch phase traces of every tap-GEMM of one VAE decode (ACE_B200_GEMM_TRACE=1) + ACE_GGML_VAE_PROFILE line."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ace-step-1.5-ggml_b200"), os.path.join(ROOT, "tests")]
from parity_common import full_vae_dir, turbo_dit_dir  # noqa: E402
from acestep_b200.bridge import B200Bridge  # noqa: E402

secs = int(sys.argv[1]) if len(sys.argv) > 1 else 60
br = B200Bridge()
br.load_vae(full_vae_dir())
lat = np.random.default_rng(0).standard_normal((secs * 25, 64), dtype=np.float32)
for _ in range(2):
    br.vae_decode_tfirst(lat)
    print(f"decode {secs} s: device {br.last_device_ms():.2f} ms", flush=True)
os.environ["ACE_B200_GEMM_TRACE"] = "1"
br.vae_decode_tfirst(lat)
br.close()
