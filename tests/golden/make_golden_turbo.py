#!/usr/bin/env python
"""Regenerates tests/golden/turbo_*.npz: outputs of the UNMODIFIED reference (oracle/_ref, ggml CPU backend) at the REAL
model dimensions of BASELINE.json's configs (turbo DiT: hidden 2048, 24 layers, 16/8 heads x 128, MLP 6144; full Oobleck VAE:
2048 -> 128 channels, ratios 10/6/4/4/2).  Run in the build container (8 cores, ~20 GiB RAM, ~20 min):

    make -C oracle ref && python tests/golden/make_golden_turbo.py [case ...]

Weights and inputs are NOT stored — they are regenerated from seeds (tools/fixtures.py: torch CPU generator for the 1.6 G
parameter fixtures, ggml_quantize_chunk for the quantised GGUF twins).  `wsum_*` entries are float64 checksums of a few
quantised/raw tensors of the fixture so that a test on another machine can tell "the fixture differs" from "the kernel
differs".

  turbo_q8_30s.npz     configs[0]: Q8_0 GGUF, T=750 (30 s), E=512, inputs seed 1234
     v_layers{1,2,6,12,24}  forward at t = t_r = 0.7 with ACE_GGML_DIT_MAX_LAYERS = k   (per DiT block)
     x_steps                8-step Euler trace, shift-3 schedule                           (per denoise step)  [8, 750, 64]
     x_cfg2                 2 Euler steps (t = 1.0, 0.5) with CFG: v = v_u + 7 (v_c - v_u), v_u from all-zero encoder states,
                            composed from two reference forwards per step (the reference ABI has no CFG entry)
  turbo_f16_cfg.npz    F16 GGUF, T=750: x_cfg2 as above (F16 keeps the reference's own activation rounding at 1e-3, so the 7x
                            amplification of CFG leaves a tight bound)
  turbo_f16_60s.npz    configs[1]: F16 GGUF, T=1500 (60 s), E=512, seed 1234:  v (t = 1.0), v1 (t = 0.9)
                            (the same two evaluations bench.py's cpu_baseline leg times)
  turbo_kq_240s.npz    configs[2] shape: Q4_K / Q6_K GGUF, T=6000 (240 s, S=3000), E=512, seed 1234, k = 1, 2 layers (what the CPU's
                            RAM allows: the reference materialises [S,S,16] F32 scores).  Stored: every 4th frame (rows 0,4,8,..)
  vae_full.npz         full VAE: audio_{8,40} = decode of latents seed 7
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ace-step-1.5-ggml_b200"), os.path.join(ROOT, "tests")]

from tools import fixtures as fx  # noqa: E402
from oracle import refbridge  # noqa: E402
from parity_common import turbo_dit_dir, full_vae_dir, fixture_checksums, shift_schedule  # noqa: E402

THREADS = os.cpu_count() or 8
GIB = 1 << 30


def fwd(d, inp, t, k=None, buf=8 * GIB, hidden=None, enc=None):
    env = {"ACE_GGML_DIT_MAX_LAYERS": str(k)} if k else {}
    i = dict(inp)
    if hidden is not None:
        i["hidden_states"] = hidden
    if enc is not None:
        i["encoder_hidden_states"] = enc
    t0 = time.time()
    outs, _ = refbridge.run_ref([{"op": "dit_forward", "t": t, "t_r": t, "out": "v"}], i, dit_dir=d, threads=THREADS,
                                compute_buffer_bytes=buf, env=env, timeout=7200)
    print(f"   forward T={i['hidden_states'].shape[0]} k={k} t={t}: {time.time() - t0:.1f}s", flush=True)
    return outs["v"]


def case_q8_30s():
    d = turbo_dit_dir("Q8_0")
    inp = fx.dit_inputs(fx.TURBO, 750, 512, seed=1234)
    g = dict(fixture_checksums(d))
    for k in (1, 2, 6, 12, 24):
        g[f"v_layers{k}"] = fwd(d, inp, 0.7, k)
    sched = [float(x) for x in shift_schedule(3.0, 8)]
    t0 = time.time()
    outs, _ = refbridge.run_ref([{"op": "dit_sample", "steps": sched, "out": "x_final", "trace": "x_steps"}], inp, dit_dir=d,
                                threads=THREADS, compute_buffer_bytes=8 * GIB, timeout=7200)
    print(f"   8-step trace: {time.time() - t0:.1f}s", flush=True)
    g["x_steps"] = outs["x_steps"]
    g["x_cfg2"] = cfg2(d, inp)
    return g


def cfg2(d, inp):
    """two Euler steps with CFG composed from two reference forwards per step"""
    x = inp["hidden_states"].copy()
    zero_enc = np.zeros_like(inp["encoder_hidden_states"])
    ts = [1.0, 0.5]
    for i, t in enumerate(ts):
        vc = fwd(d, inp, t, hidden=x)
        vu = fwd(d, inp, t, hidden=x, enc=zero_enc)
        v = vu + np.float32(7.0) * (vc - vu)
        dt = t if i == len(ts) - 1 else t - ts[i + 1]
        x = (x - v * np.float32(dt)).astype(np.float32)
    return x


def case_f16_cfg():
    d = turbo_dit_dir("F16")
    inp = fx.dit_inputs(fx.TURBO, 750, 512, seed=1234)
    g = dict(fixture_checksums(d))
    g["x_cfg2"] = cfg2(d, inp)
    return g


def case_f16_60s():
    d = turbo_dit_dir("F16")
    inp = fx.dit_inputs(fx.TURBO, 1500, 512, seed=1234)
    g = dict(fixture_checksums(d))
    g["v"] = fwd(d, inp, 1.0, buf=20 * GIB)
    g["v1"] = fwd(d, inp, 0.9, buf=20 * GIB)
    return g


def case_kq_240s():
    g = {}
    inp = fx.dit_inputs(fx.TURBO, 6000, 512, seed=1234)
    for q in ("Q4_K", "Q6_K"):
        d = turbo_dit_dir(q)
        for k2, v in fixture_checksums(d).items():
            g[f"{k2}_{q}"] = v
        for k in (1, 2):
            g[f"v_{q}_layers{k}"] = fwd(d, inp, 0.7, k, buf=(10 + 9 * k) * GIB)[::4].copy()
    return g


def case_vae_full():
    d = full_vae_dir()
    g = {}
    for n in (8, 40):
        lat = fx.vae_latents(fx.VAE_FULL, n, seed=7)
        t0 = time.time()
        o, _ = refbridge.run_ref([{"op": "vae_decode", "latents": "lat", "out": "audio"}], {"lat": lat}, vae_dir=d, threads=THREADS,
                                 compute_buffer_bytes=(2 + n // 2) * GIB, timeout=7200)
        print(f"   vae decode {n} frames: {time.time() - t0:.1f}s", flush=True)
        g[f"audio_{n}"] = o["audio"]
    return g


CASES = {"turbo_q8_30s": case_q8_30s, "turbo_f16_60s": case_f16_60s, "turbo_f16_cfg": case_f16_cfg, "turbo_kq_240s": case_kq_240s, "vae_full": case_vae_full}


def main():
    assert refbridge.ref_available(), "build the reference first: make -C oracle ref"
    for name in (sys.argv[1:] or list(CASES)):
        print(f"[{name}]", flush=True)
        g = CASES[name]()
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **{k: (v.astype(np.float32) if np.asarray(v).dtype == np.float32 else np.asarray(v)) for k, v in g.items()})
        print(f"  wrote {path}: {os.path.getsize(path)} bytes", flush=True)


if __name__ == "__main__":
    main()
