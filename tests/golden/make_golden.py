#!/usr/bin/env python
"""Regenerates tests/golden/*.npz: outputs of the UNMODIFIED reference (oracle/_ref, built from /root/reference by
oracle/Makefile, ggml CPU backend) on the seeded toy fixtures.  Run in the build container:

    make -C oracle ref && python tests/golden/make_golden.py

Inputs and weights are NOT stored: they are regenerated from seeds by acestep_b200.fixtures (numpy Generator,
stable across versions), so the files hold only reference outputs (a few hundred KB in total).

  dit_toy.npz   TOY dims (hidden 256, 4 layers, 2/1 heads x 128, window 8), weights seed 0
     v_f32            full forward, F32 safetensors, T=33 E=7 inputs seed 1234, t=t_r=0.6
     v_f32_layers{k}  same network truncated to k=1..4 layers (ACE_GGML_DIT_MAX_LAYERS), T=40 E=9 seed 77, t=0.3
     v_masked         T=48 E=12 seed 5, attention_mask[40:]=0, encoder_attention_mask[9:]=0, t=0.8 t_r=0.5
     v_noenc          T=48 E=0 seed 6, t=0.8
     x_steps          8-step Euler trace (shift 3), T=32 E=8 seed 11           [8, 32, 64]
     v_{F16,Q8_0,Q4_K,Q6_K}   full forward on the GGUF twin of the same weights, T=33 E=7 seed 1234, t=0.6
  vae_toy.npz   VAE_TOY dims (channels 128->64->32->32, ratios [6,4,2] up), weights seed 0
     audio_{1,16,45}  decode of latents seed 7 with n_frames frames
     audio_tiled_60   chunk 36 / overlap 12 tiled decode of 60 frames (seed 9)
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ace-step-1.5-ggml_b200"), os.path.join(ROOT, "tests")]

from tools import fixtures as fx  # noqa: E402
from oracle import refbridge  # noqa: E402
from parity_common import make_dit_dir, make_vae_dir, ref_dit_forward  # noqa: E402


def main():
    assert refbridge.ref_available(), "build the reference first: make -C oracle ref"
    root = tempfile.mkdtemp(prefix="ace_golden_")
    dit = {}
    d = make_dit_dir(root, fx.TOY, 0, None)
    inp = fx.dit_inputs(fx.TOY, 33, 7, seed=1234)
    dit["v_f32"], _ = ref_dit_forward(d, inp, 0.6, 0.6)
    inp = fx.dit_inputs(fx.TOY, 40, 9, seed=77)
    for k in (1, 2, 3, 4):
        dit[f"v_f32_layers{k}"], _ = ref_dit_forward(d, inp, 0.3, 0.3, max_layers=k)
    inp = fx.dit_inputs(fx.TOY, 48, 12, seed=5)
    am = np.ones((48,), np.int32)
    am[40:] = 0
    em = np.ones((12,), np.int32)
    em[9:] = 0
    dit["v_masked"], _ = ref_dit_forward(d, inp, 0.8, 0.5, masks={"mask": am, "enc_mask": em})
    inp = fx.dit_inputs(fx.TOY, 48, 0, seed=6)
    dit["v_noenc"], _ = ref_dit_forward(d, inp, 0.8, 0.8)
    inp = fx.dit_inputs(fx.TOY, 32, 8, seed=11)
    u = 1.0 - np.arange(8, dtype=np.float64) / 8
    sched = [float(x) for x in (3.0 * u / (1.0 + 2.0 * u)).astype(np.float32)]
    outs, _ = refbridge.run_ref([{"op": "dit_sample", "steps": sched, "out": "x_final", "trace": "x_steps"}], inp, dit_dir=d)
    dit["x_steps"] = outs["x_steps"]
    inp = fx.dit_inputs(fx.TOY, 33, 7, seed=1234)
    for q in ("F16", "Q8_0", "Q4_K", "Q6_K"):
        dq = make_dit_dir(root, fx.TOY, 0, q)
        dit[f"v_{q}"], _ = ref_dit_forward(dq, inp, 0.6, 0.6)
    np.savez_compressed(os.path.join(HERE, "dit_toy.npz"), **{k: v.astype(np.float32) for k, v in dit.items()})

    vae = {}
    vd = make_vae_dir(root)
    for n in (1, 16, 45):
        lat = fx.vae_latents(fx.VAE_TOY, n, seed=7)
        o, _ = refbridge.run_ref([{"op": "vae_decode", "latents": "lat", "out": "audio"}], {"lat": lat}, vae_dir=vd)
        vae[f"audio_{n}"] = o["audio"]
    lat = fx.vae_latents(fx.VAE_TOY, 60, seed=9)
    o, _ = refbridge.run_ref([{"op": "vae_decode_tiled", "latents": "lat", "out": "audio", "chunk": 36, "overlap": 12}], {"lat": lat}, vae_dir=vd)
    vae["audio_tiled_60"] = o["audio"]
    np.savez_compressed(os.path.join(HERE, "vae_toy.npz"), **{k: v.astype(np.float32) for k, v in vae.items()})
    for f in ("dit_toy.npz", "vae_toy.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")


if __name__ == "__main__":
    main()
