#!/usr/bin/env python
"""Exploratory timing at the real model dimensions (not the bench contract; see bench.py for that).
usage: python scripts/perf_probe.py [--quant F16] [--seconds 60] [--batch 1] [--steps 8] [--vae-seconds 30] [--cpu]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ace-step-1.5-ggml_b200")]

from tools import fixtures as fx  # noqa: E402
from acestep_b200.bridge import B200Bridge  # noqa: E402


def fixture_dir(kind, quant=None):
    root = os.environ.get("ACE_B200_FIXTURES", "/tmp/ace_b200_fixtures")
    os.makedirs(root, exist_ok=True)
    if kind == "dit":
        d = os.path.join(root, f"turbo_{quant}")
        if not os.path.exists(os.path.join(d, "model.gguf")):
            t = time.time()
            lib = None
            if quant != "F16":
                from oracle import refbridge
                lib = refbridge.ref_lib()
            fx.make_dit_fixture(d, fx.TURBO, 0, safetensors=False, gguf_quant=quant, ggml_lib=lib, fast=True)
            print(f"[fixture] wrote {d} in {time.time() - t:.1f}s", flush=True)
        return d
    d = os.path.join(root, "vae_full")
    if not os.path.exists(os.path.join(d, "config.json")):
        t = time.time()
        fx.make_vae_fixture(d, fx.VAE_FULL, 0, fast=True)
        print(f"[fixture] wrote {d} in {time.time() - t:.1f}s", flush=True)
    return d


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quant", default="F16")
    ap.add_argument("--seconds", type=float, default=60)
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--enc", type=int, default=512)
    ap.add_argument("--cfg", action="store_true")
    ap.add_argument("--vae-seconds", type=float, default=0)
    ap.add_argument("--cpu", action="store_true", help="also time one reference CPU evaluation")
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()

    br = B200Bridge()
    T = int(round(a.seconds * 25))
    if a.seconds > 0:
        d = fixture_dir("dit", a.quant)
        t = time.time()
        br.load_dit(d)
        print(f"[load] dit {time.time() - t:.1f}s  {br.model_info()}", flush=True)
        g = np.random.default_rng(1234)
        B, E = a.batch, a.enc
        noise = g.standard_normal((B, T, 64), dtype=np.float32)
        ctx = np.ones((B, T, 128), dtype=np.float32)
        ctx[:, :, :64] = g.standard_normal((B, T, 64), dtype=np.float32)
        enc = g.standard_normal((B, E, 2048), dtype=np.float32)
        un = np.zeros_like(enc) if a.cfg else None
        sched = br.shift_schedule(3.0, a.steps)
        for r in range(a.reps):
            n0 = br.kernel_launches()
            t = time.time()
            lat = br.generate_latents(noise, ctx, enc, un, schedule=sched, guidance=7.0)
            wall = time.time() - t
            ms = br.last_device_ms()
            print(f"[dit] rep{r} T={T} B={B} cfg={a.cfg} steps={a.steps}: device {ms:.2f} ms ({ms / a.steps:.3f} ms/step), wall {wall * 1e3:.1f} ms, "
                  f"launches {br.kernel_launches() - n0}, finite={np.isfinite(lat).all()} rms={float(np.sqrt((lat ** 2).mean())):.4f}", flush=True)
        br.profile_enable(True)
        br.generate_latents(noise, ctx, enc, un, schedule=sched, guidance=7.0)
        prof = br.profile_collect()
        br.profile_enable(False)
        for k, v in prof.items():
            if v["launches"]:
                tf = v["flops"] / (v["ms"] * 1e-3) / 1e12 if v["ms"] > 0 else 0
                gb = v["bytes"] / (v["ms"] * 1e-3) / 1e9 if v["ms"] > 0 else 0
                print(f"[prof] {k:12s} {v['ms']:9.3f} ms  {v['launches']:6d} launches  {tf:8.1f} TFLOP/s  {gb:8.1f} GB/s(alg)", flush=True)
        # single-call ABI (host round trip every step)
        t = time.time()
        v = br.dit_forward_tfirst(noise[0], ctx[0], enc[0], None, None, 1.0, 1.0)
        print(f"[abi] one ace_ggml_dit_forward: wall {(time.time() - t) * 1e3:.1f} ms, device {br.last_device_ms():.2f} ms", flush=True)
        if a.cpu:
            from oracle import refbridge
            inp = {"hidden_states": noise[0], "context_latents": ctx[0], "encoder_hidden_states": enc[0]}
            ncpu = os.cpu_count() or 1
            outs, tim = refbridge.run_ref([{"op": "dit_forward", "t": 1.0, "t_r": 1.0, "out": "v"}], inp, dit_dir=d, threads=ncpu,
                                          compute_buffer_bytes=int(os.environ.get("ACE_REF_BUFFER_GB", "24")) << 30,
                                          env={"ACE_GGML_DIT_PROFILE": "1"})
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            from parity_common import metrics
            print(f"[cpu] reference one evaluation: {tim['dit_forward_v_s']:.2f}s on {ncpu} threads; load {tim['load_s']:.1f}s", flush=True)
            print("[cpu] stderr:", tim["stderr"].strip().splitlines()[-1:], flush=True)
            print("[parity] vs reference:", json.dumps(metrics(v, outs["v"])), flush=True)
    if a.vae_seconds > 0:
        d = fixture_dir("vae")
        t = time.time()
        br.load_vae(d)
        print(f"[load] vae {time.time() - t:.1f}s  {br.model_info()}", flush=True)
        Tv = int(round(a.vae_seconds * 25))
        lat = fx.vae_latents(fx.VAE_FULL, Tv, seed=7)
        for r in range(a.reps):
            t = time.time()
            au = br.vae_decode_tfirst(lat)
            print(f"[vae] rep{r} frames={Tv}: device {br.last_device_ms():.2f} ms, wall {(time.time() - t) * 1e3:.1f} ms, finite={np.isfinite(au).all()} "
                  f"rms={float(np.sqrt((au ** 2).mean())):.4f}", flush=True)
        br.profile_enable(True)
        br.vae_decode_tfirst(lat)
        prof = br.profile_collect()
        br.profile_enable(False)
        for k, v in prof.items():
            if v["launches"]:
                tf = v["flops"] / (v["ms"] * 1e-3) / 1e12 if v["ms"] > 0 else 0
                print(f"[prof] {k:12s} {v['ms']:9.3f} ms  {v['launches']:6d} launches  {tf:8.1f} TFLOP/s", flush=True)
    br.close()


if __name__ == "__main__":
    main()
