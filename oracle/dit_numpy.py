"""TEST INFRASTRUCTURE (oracle). NumPy restatement of the reference's DiT velocity network and sampler.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module; the product (libacestep_b200.so) never routes through it.

What it restates (file:line in /root/reference/acestep_ggml/cpp unless noted):
  forward_dit            acestep_dit_model.cpp:1316-1560
  attention              acestep_dit_model.cpp:1175-1259   (mask rule :1132-1173, GQA repeat :1118-1130)
  timestep_forward       acestep_dit_model.cpp:1261-1308
  rms_norm               acestep_dit_model.cpp:1097-1106 + ggml_rms_norm (ggml-cpu/ops.cpp:3646-3697: sum in double)
  NEOX RoPE              ggml_rope_ext, ggml-cpu/ops.cpp:5637-5777 (pairs (i, i+hd/2), angle pos*theta^(-2i/hd))
  weight relayout        acestep_dit_model.cpp:334-490  (proj_in [out,c,k] -> [(k*in_ch+c), out]; proj_out likewise)
  Euler update/schedule  acestep_ggml.cpp:1484-1500, 2056-2086
  CFG combine            v = v_u + g (v_c - v_u)  (by composition; the reference ABI has no CFG: SURVEY.md 8c)

Arithmetic is float64 with float32 inputs: this is the "F32 truth" of SURVEY.md 7.3(1) — the reference's F32-weight
path (ggml f32 dot products) agrees with it to ~1e-6 relative, which tests/test_oracle_cpu.py pins against golden
vectors produced by the UNMODIFIED reference built from source (oracle/_ref; tests/golden/make_golden.py).
Quantised-weight parity is pinned through the reference itself (oracle/_ref), not through this file.
"""
from __future__ import annotations

from typing import Dict, Optional, Sequence

import numpy as np

F = np.float64


def _lin(x, w, b=None):
    y = x @ w.astype(F).T
    return y if b is None else y + b.astype(F)


def rms_norm(x, w, eps):
    """ggml_rms_norm over the last axis, then * weight"""
    ms = np.mean(x * x, axis=-1, keepdims=True)
    return x / np.sqrt(ms + eps) * w.astype(F)


def silu(x):
    return x / (1.0 + np.exp(-x))


def timestep_features(t: float, dim: int = 256, scale: float = 1000.0):
    """build_timestep_freq: [cos(0..half) | sin(0..half)] of t*scale*exp(-ln(1e4) i/half), computed in f32 like the reference"""
    half = dim // 2
    i = np.arange(half, dtype=np.float32)
    f = np.exp(-np.float32(np.log(np.float32(10000.0))) * i / np.float32(half)).astype(np.float32)
    arg = (np.float32(t) * np.float32(scale)) * f
    return np.concatenate([np.cos(arg), np.sin(arg)]).astype(F)


def timestep_forward(T: Dict[str, np.ndarray], prefix: str, t: float):
    x = timestep_features(t)
    temb = _lin(silu(_lin(x, T[prefix + ".linear_1.weight"], T[prefix + ".linear_1.bias"])),
                T[prefix + ".linear_2.weight"], T[prefix + ".linear_2.bias"])
    proj = _lin(silu(temb), T[prefix + ".time_proj.weight"], T[prefix + ".time_proj.bias"])
    return temb, proj.reshape(6, -1)


def rope_neox(x, theta: float):
    """x: [S, heads, hd]; position = row index"""
    S, _, hd = x.shape
    half = hd // 2
    inv = np.float64(theta) ** (-2.0 * np.arange(half, dtype=F) / hd)
    ang = np.arange(S, dtype=F)[:, None] * inv[None, :]
    c, s = np.cos(ang)[:, None, :], np.sin(ang)[:, None, :]
    a, b = x[..., :half], x[..., half:]
    return np.concatenate([a * c - b * s, a * s + b * c], axis=-1)


def attention(T, cfg, prefix, q_in, kv_in, key_mask, sliding: bool, use_rope: bool):
    nq, nkv, hd = cfg["num_attention_heads"], cfg["num_key_value_heads"], cfg["head_dim"]
    eps = cfg["rms_norm_eps"]
    S, Kl = q_in.shape[0], kv_in.shape[0]
    q = _lin(q_in, T[prefix + ".q_proj.weight"]).reshape(S, nq, hd)
    k = _lin(kv_in, T[prefix + ".k_proj.weight"]).reshape(Kl, nkv, hd)
    v = _lin(kv_in, T[prefix + ".v_proj.weight"]).reshape(Kl, nkv, hd)
    q = rms_norm(q, T[prefix + ".q_norm.weight"], eps)
    k = rms_norm(k, T[prefix + ".k_norm.weight"], eps)
    if use_rope:
        q = rope_neox(q, cfg["rope_theta"])
        k = rope_neox(k, cfg["rope_theta"])
    rep = nq // nkv
    k = np.repeat(k, rep, axis=1)  # head h uses kv head h // rep (repeat_kv_interleave)
    v = np.repeat(v, rep, axis=1)
    sc = np.einsum("qhd,khd->hqk", q, k) / np.sqrt(np.float32(hd)).astype(F)
    allow = np.ones((S, Kl), dtype=bool)
    if sliding:
        d = np.arange(S)[:, None] - np.arange(Kl)[None, :]
        allow &= np.abs(d) <= cfg["sliding_window"]
    if key_mask is not None:
        allow &= (np.asarray(key_mask) != 0)[None, :]
    sc = np.where(allow[None], sc, -np.inf)
    m = sc.max(axis=-1, keepdims=True)
    p = np.exp(sc - m)
    p /= p.sum(axis=-1, keepdims=True)
    o = np.einsum("hqk,khd->qhd", p, v).reshape(S, nq * hd)
    return _lin(o, T[prefix + ".o_proj.weight"])


def forward_dit(T: Dict[str, np.ndarray], cfg: dict, hidden_states, context_latents, encoder_hidden_states=None,
                attention_mask=None, encoder_attention_mask=None, timestep: float = 1.0, timestep_r: Optional[float] = None,
                max_layers: Optional[int] = None, trace: Optional[list] = None):
    """hidden_states [Tn, A], context_latents [Tn, ctx], encoder_hidden_states [E, H] -> velocity [Tn, A] (float64).
    `trace` (a list) receives the residual stream after every layer."""
    if timestep_r is None:
        timestep_r = timestep
    A, P, C, H = cfg["audio_acoustic_hidden_dim"], cfg["patch_size"], cfg["in_channels"], cfg["hidden_size"]
    ctx = C - A
    eps = cfg["rms_norm_eps"]
    Tn = hidden_states.shape[0]
    S = (Tn + P - 1) // P
    x0 = np.zeros((S * P, C), dtype=F)
    if context_latents is not None:
        x0[:Tn, :ctx] = context_latents
    x0[:Tn, ctx:] = hidden_states
    # proj_in: W2[out, k*C + c] = w[out, c, k]
    w_in = T["decoder.proj_in.1.weight"].astype(F).transpose(0, 2, 1).reshape(H, P * C)
    x = x0.reshape(S, P * C) @ w_in.T + T["decoder.proj_in.1.bias"].astype(F)

    enc = None
    E = 0 if encoder_hidden_states is None else encoder_hidden_states.shape[0]
    if E > 0:
        enc = _lin(encoder_hidden_states.astype(F), T["decoder.condition_embedder.weight"], T["decoder.condition_embedder.bias"])

    temb_t, proj_t = timestep_forward(T, "decoder.time_embed", timestep)
    temb_r, proj_r = timestep_forward(T, "decoder.time_embed_r", np.float32(timestep) - np.float32(timestep_r))
    temb, proj = temb_t + temb_r, proj_t + proj_r

    pmask = None
    if attention_mask is not None:
        am = np.zeros((S * P,), dtype=np.int32)
        am[:Tn] = np.asarray(attention_mask)[:Tn]
        pmask = (am.reshape(S, P) != 0).any(axis=1).astype(np.int32)

    n_layers = cfg["num_hidden_layers"] if not max_layers else min(max_layers, cfg["num_hidden_layers"])
    for i in range(n_layers):
        p = f"decoder.layers.{i}."
        ss = T[p + "scale_shift_table"].astype(F).reshape(6, H) + proj
        shift_msa, scale_msa, gate_msa, c_shift, c_scale, c_gate = ss
        n = rms_norm(x, T[p + "self_attn_norm.weight"], eps) * (1.0 + scale_msa) + shift_msa
        sliding = cfg["layer_types"][i] == "sliding_attention" and cfg.get("use_sliding_window", True)
        x = x + attention(T, cfg, p + "self_attn", n, n, pmask, sliding, True) * gate_msa
        if enc is not None:
            cn = rms_norm(x, T[p + "cross_attn_norm.weight"], eps)
            x = x + attention(T, cfg, p + "cross_attn", cn, enc, encoder_attention_mask, False, False)
        mn = rms_norm(x, T[p + "mlp_norm.weight"], eps) * (1.0 + c_scale) + c_shift
        act = silu(_lin(mn, T[p + "mlp.gate_proj.weight"])) * _lin(mn, T[p + "mlp.up_proj.weight"])
        x = x + _lin(act, T[p + "mlp.down_proj.weight"]) * c_gate
        if trace is not None:
            trace.append(x.copy())

    oss = T["decoder.scale_shift_table"].astype(F).reshape(2, H) + temb
    y = rms_norm(x, T["decoder.norm_out.weight"], eps) * (1.0 + oss[1]) + oss[0]
    # proj_out: y_lin[p, k*A + a] = sum_h y[p, h] w[h, a, k]
    w_out = T["decoder.proj_out.1.weight"].astype(F).transpose(2, 1, 0).reshape(P * A, H)
    out = (y @ w_out.T).reshape(S * P, A) + T["decoder.proj_out.1.bias"].astype(F)
    return out[:Tn]


def shift_schedule(shift: float, n_steps: int) -> np.ndarray:
    """t_i = s u / (1 + (s-1) u), u = 1 - i/n: closed form of the three 8-step tables at acestep_ggml.cpp:1485-1487"""
    u = 1.0 - np.arange(n_steps, dtype=np.float64) / n_steps
    return (shift * u / (1.0 + (shift - 1.0) * u)).astype(np.float32)


def euler_sample(T, cfg, noise, context_latents, encoder_hidden_states, schedule: Sequence[float],
                 uncond_encoder_hidden_states=None, guidance: float = 1.0, attention_mask=None, encoder_attention_mask=None):
    """x <- x - v (t_i - t_{i+1}); last step x <- x - v t_i  (acestep_ggml.cpp:2056-2086). Returns (x_final, [x_i])."""
    x = noise.astype(np.float32)
    steps = []
    for i, t in enumerate(schedule):
        t = float(t)
        v = forward_dit(T, cfg, x, context_latents, encoder_hidden_states, attention_mask, encoder_attention_mask, t, t)
        if uncond_encoder_hidden_states is not None:
            vu = forward_dit(T, cfg, x, context_latents, uncond_encoder_hidden_states, attention_mask, encoder_attention_mask, t, t)
            v = vu + guidance * (v - vu)
        dt = np.float32(t) if i == len(schedule) - 1 else np.float32(t) - np.float32(schedule[i + 1])
        x = (x - v.astype(np.float32) * dt).astype(np.float32)
        steps.append(x.copy())
    return x, steps
