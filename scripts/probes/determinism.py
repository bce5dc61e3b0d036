#!/usr/bin/env python
"""Run-to-run determinism of full DiT evaluations (turbo dims, T = 6000): N identical calls, how many distinct outputs?

    python scripts/probes/determinism.py [reps]            one evaluation (M = 3000) under the route switches, Q4_K / F16 / Q8_0
    python scripts/probes/determinism.py [reps] sizes      batches of 1 / 2 / 16 evaluations (M = 3000 / 6000 / 48 000), Q4_K / Q6_K / Q8_0
    python scripts/probes/determinism.py [reps] pub        the publish mode of the pair kernel (ACE_B200_GEMM_QPUB=1)
    python scripts/probes/determinism.py [reps] fix        in-kernel decode (ACE_B200_DECODE_PASS=0): cluster-scoped decoder fence vs the
                                                           CTA-scoped one (ACE_B200_GEMM_DBG=1) that let a tensor core read a stale sub-tile

This is the probe that found and bisected the visibility race of the CTA-pair kernel's in-kernel decoders (DESIGN.md 4.1,
profiles/r2_determinism_investigation.log; the investigation build had more ACE_B200_GEMM_DBG bits — fence placements in the relay / the
leader / one lane per warp — than the library keeps)."""
import hashlib
import os
import sys
from collections import Counter

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ace-step-1.5-ggml_b200"), os.path.join(ROOT, "tests")]
from tools import fixtures as fx  # noqa: E402
from acestep_b200.bridge import B200Bridge  # noqa: E402
from parity_common import metrics, turbo_dit_dir  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
T = 6000
inp = fx.dit_inputs(fx.TURBO, T, 512, seed=1236)


def trial(quant, env, label):
    for k in [k for k in os.environ if k.startswith("ACE_B200_") and k != "ACE_B200_DEVICE"]:
        del os.environ[k]
    os.environ.update(env)
    br = B200Bridge()
    br.load_dit(turbo_dit_dir(quant))
    outs, hashes = {}, []
    for _ in range(reps):
        v = br.dit_forward_tfirst(inp["hidden_states"], inp["context_latents"], inp["encoder_hidden_states"], None, None, 0.3, 0.3)
        hsh = hashlib.md5(v.tobytes()).hexdigest()
        hashes.append(hsh)
        outs.setdefault(hsh, v)
    cnt = Counter(hashes)
    mode = cnt.most_common(1)[0][0]
    worst = max((metrics(v, outs[mode])["rel_l2"] for h, v in outs.items() if h != mode), default=0.0)
    print(f"[{quant:5s}] {label:34s} {len(cnt):3d} distinct outputs in {reps} calls (mode x{cnt[mode]}), worst rel_l2 to the mode {worst:.2e}", flush=True)
    br.close()


def trial_batched(quant, B, n, env, label):
    for k in [k for k in os.environ if k.startswith("ACE_B200_") and k != "ACE_B200_DEVICE"]:
        del os.environ[k]
    os.environ.update(env)
    br = B200Bridge()
    br.load_dit(turbo_dit_dir(quant))
    ins = [fx.dit_inputs(fx.TURBO, T, 512, seed=1234 + i) for i in range(B)]
    h = np.stack([i["hidden_states"] for i in ins]); c = np.stack([i["context_latents"] for i in ins]); e = np.stack([i["encoder_hidden_states"] for i in ins])
    ts = [0.3 + 0.04 * i for i in range(B)]
    outs, hashes = {}, []
    for _ in range(n):
        v = br.dit_forward_batched(h, c, e, None, None, ts, ts)
        hsh = hashlib.md5(v.tobytes()).hexdigest()
        hashes.append(hsh)
        outs.setdefault(hsh, v.copy())
    cnt = Counter(hashes)
    mode = cnt.most_common(1)[0][0]
    worst = max((metrics(v, outs[mode])["rel_l2"] for hh, v in outs.items() if hh != mode), default=0.0)
    print(f"[{quant:5s}] {label:28s} batch {B:2d} (M = {B * 3000:5d}): {len(cnt):3d} distinct outputs in {n} calls = {n * B} evaluations, "
          f"worst rel_l2 to the mode {worst:.2e}", flush=True)
    br.close()


if len(sys.argv) > 2 and sys.argv[2] == "pub":
    for q in ("Q4_K", "Q6_K", "Q8_0"):
        env = {"ACE_B200_GEMM_QPUB": "1", "ACE_B200_DECODE_PASS": "0"}
        trial_batched(q, 16, max(4, reps // 12), env, "publish mode")
        trial_batched(q, 1, reps // 2, env, "publish mode")
    sys.exit(0)
if len(sys.argv) > 2 and sys.argv[2] == "sizes":
    for q in ("Q4_K", "Q6_K", "Q8_0"):
        trial_batched(q, 1, reps, {}, "default")
        trial_batched(q, 2, reps // 2, {}, "default")
    trial_batched("Q4_K", 16, max(4, reps // 12), {}, "default")
    sys.exit(0)
if len(sys.argv) > 2 and sys.argv[2] == "fix":
    trial("Q4_K", {"ACE_B200_DECODE_PASS": "0"}, "in-kernel decode, cluster-scoped fence")
    trial("Q4_K", {"ACE_B200_DECODE_PASS": "0", "ACE_B200_GEMM_DBG": "1"}, "in-kernel decode, CTA-scoped fence (racy)")
    trial_batched("Q4_K", 16, max(4, reps // 12), {"ACE_B200_DECODE_PASS": "0"}, "in-kernel, cluster fence")
    trial_batched("Q4_K", 16, max(4, reps // 12), {"ACE_B200_DECODE_PASS": "0", "ACE_B200_GEMM_DBG": "1"}, "in-kernel, CTA fence (racy)")
    sys.exit(0)
trial("Q4_K", {}, "default (decode pass from 2048 rows)")
trial("Q4_K", {"ACE_B200_DECODE_PASS": "0"}, "in-kernel decode everywhere")
trial("Q4_K", {"ACE_B200_DECODE_PASS": "0", "ACE_B200_GEMM_2CTA": "0"}, "single-CTA GEMMs, in-kernel decode")
trial("Q4_K", {"ACE_B200_PREDECODE": "1"}, "F16 twin through the pair kernel")
trial("Q4_K", {"ACE_B200_ATTN": "tc4"}, "attention: smem-P kernel")
trial("Q4_K", {"ACE_B200_ATTN": "mma"}, "attention: mma.sync kernel")
trial("Q4_K", {"ACE_B200_GEMM_TMARED": "0"}, "red.add instead of TMA reduce")
trial("Q4_K", {"ACE_B200_GEMM_WARM": "0"}, "no epilogue warm-up pass")
trial("Q4_K", {"ACE_B200_GEMM_QDEC": "1"}, "one decoder group")
trial("Q4_K", {"ACE_B200_GEMM_QEPI": "4"}, "4 epilogue warps everywhere")
trial("Q4_K", {"ACE_B200_GEMM_BAND_MB": "0"}, "no rasterisation bands")
trial("Q4_K", {"ACE_B200_NO_SCHED_TEMB": "1", "ACE_B200_NO_COND_CACHE": "1"}, "no conditioning cache")
trial("F16", {}, "F16 weights, default")
trial("F16", {"ACE_B200_GEMM_2CTA": "0"}, "F16 weights, single-CTA GEMMs")
trial("Q8_0", {}, "Q8_0 weights, default")
