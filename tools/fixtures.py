"""Seeded random-init weight fixtures for the DiT and the Oobleck VAE decoder.

There are no checkpoints in the reference tree, so every parity test and the
benchmark run on random-init weights written in the exact on-disk layouts the
reference loaders accept:

* DiT: ``config.json`` + ``model.safetensors`` with the tensor names/shapes of
  ``acestep_ggml/cpp/acestep_dit_model.cpp:870-1082`` (PyTorch shapes), and GGUF
  twins following the exporter's rule
  (``acestep_ggml/tools/export_safetensors_to_gguf.py:213-265``: quantise every
  >=2-D float tensor whose last dim is a multiple of the block size, everything
  else F16).
* VAE: ``config.json`` + ``diffusion_pytorch_model.safetensors`` with weight-norm
  ``weight_g``/``weight_v`` pairs (``acestep_vae_model.cpp:520-588, 761-955``);
  the reference loader also insists on the ``encoder.*`` tensors.

Nothing here touches the GPU or the oracle, except that K-quant GGUF export
needs ``ggml_quantize_chunk`` which (like the reference exporter, :120-151) is
borrowed from a ggml shared library when the caller passes one.
"""
from __future__ import annotations

import ctypes
import json
import os
from dataclasses import dataclass, field, asdict
from typing import Dict, List, Optional

import numpy as np


# --------------------------------------------------------------------------- configs

@dataclass
class DitDims:
    hidden_size: int = 2048
    intermediate_size: int = 6144
    num_hidden_layers: int = 24
    num_attention_heads: int = 16
    num_key_value_heads: int = 8
    head_dim: int = 128
    max_position_embeddings: int = 32768
    rms_norm_eps: float = 1e-6
    rope_theta: float = 1000000.0
    patch_size: int = 2
    in_channels: int = 192
    audio_acoustic_hidden_dim: int = 64
    use_sliding_window: bool = True
    sliding_window: int = 128
    # condition encoders (`encoder.*` tensors, acestep_dit_model.cpp:885-994); 0 layers = not written
    text_hidden_dim: int = 1024
    num_lyric_encoder_hidden_layers: int = 0
    timbre_hidden_dim: int = 64
    num_timbre_encoder_hidden_layers: int = 0

    def layer_types(self) -> List[str]:
        # even layers sliding, odd layers full (R/acestep/mlx_dit/model.py:421-425)
        return ["sliding_attention" if (i % 2 == 0) else "full_attention"
                for i in range(self.num_hidden_layers)]

    def config_json(self) -> dict:
        d = asdict(self)
        d["layer_types"] = self.layer_types()
        d["dtype"] = "float32"
        return d


TURBO = DitDims()
# toy dims keep head_dim 128 (the attention kernel is specialised on it) and every
# GEMM K a multiple of 256 so K-quants apply to the same tensors as at turbo dims.
TOY = DitDims(hidden_size=256, intermediate_size=512, num_hidden_layers=4,
              num_attention_heads=2, num_key_value_heads=1, sliding_window=8)
MID = DitDims(hidden_size=512, intermediate_size=1536, num_hidden_layers=4,
              num_attention_heads=4, num_key_value_heads=2, sliding_window=16)
# toy DiT that also carries the condition side: text projector, 2-layer lyric encoder, 2-layer timbre encoder
TOY_COND = DitDims(hidden_size=256, intermediate_size=512, num_hidden_layers=4,
                   num_attention_heads=2, num_key_value_heads=1, sliding_window=8,
                   text_hidden_dim=256, num_lyric_encoder_hidden_layers=2, num_timbre_encoder_hidden_layers=2)


@dataclass
class QwenDims:
    """Qwen3 text encoder (qwen_config.cpp:52-62)."""
    vocab_size: int = 151669
    hidden_size: int = 1024
    num_hidden_layers: int = 28
    num_attention_heads: int = 16
    num_key_value_heads: int = 8
    intermediate_size: int = 3072
    head_dim: int = 128
    max_position_embeddings: int = 32768
    rms_norm_eps: float = 1e-6
    rope_theta: float = 1000000.0

    def config_json(self) -> dict:
        d = asdict(self)
        d["dtype"] = "float32"
        return d


QWEN_FULL = QwenDims()
QWEN_TOY = QwenDims(vocab_size=512, hidden_size=256, num_hidden_layers=3, num_attention_heads=2,
                    num_key_value_heads=1, intermediate_size=512)


@dataclass
class VaeDims:
    audio_channels: int = 2
    encoder_hidden_size: int = 128
    decoder_channels: int = 128
    decoder_input_channels: int = 64
    sampling_rate: int = 48000
    downsampling_ratios: List[int] = field(default_factory=lambda: [2, 4, 4, 6, 10])
    channel_multiples: List[int] = field(default_factory=lambda: [1, 2, 4, 8, 16])

    @property
    def hop(self) -> int:
        h = 1
        for r in self.downsampling_ratios:
            h *= r
        return h

    def config_json(self) -> dict:
        return asdict(self)


VAE_FULL = VaeDims()
# toy VAE: channels 128 -> 64 -> 32 (every level a multiple of 32, like the real 2048..128 plan)
VAE_TOY = VaeDims(decoder_channels=32, encoder_hidden_size=128,
                  downsampling_ratios=[2, 4, 6], channel_multiples=[1, 2, 4])


# --------------------------------------------------------------------------- safetensors (own writer)

_ST_DTYPES = {np.dtype(np.float32): "F32", np.dtype(np.float16): "F16"}


def save_safetensors(path: str, tensors: Dict[str, np.ndarray]) -> None:
    """8-byte LE header length + JSON header + raw little-endian data
    (the format ``acestep_ggml/cpp/safetensors.cpp:46-129`` parses)."""
    header = {}
    off = 0
    order = list(tensors.keys())
    for name in order:
        a = np.ascontiguousarray(tensors[name])
        nb = a.nbytes
        header[name] = {"dtype": _ST_DTYPES[a.dtype], "shape": list(a.shape),
                        "data_offsets": [off, off + nb]}
        off += nb
    hj = json.dumps(header, separators=(",", ":")).encode()
    pad = (-len(hj)) % 8
    hj += b" " * pad
    with open(path, "wb") as f:
        f.write(np.uint64(len(hj)).tobytes())
        f.write(hj)
        for name in order:
            f.write(np.ascontiguousarray(tensors[name]).tobytes())


# --------------------------------------------------------------------------- DiT weights

class _Rng:
    """numpy Generator for small fixtures (stable across versions); torch for the
    1.6 G-parameter turbo fixture where numpy would take minutes."""

    def __init__(self, seed: int, fast: bool):
        self.fast = fast
        if fast:
            import torch
            self.g = torch.Generator().manual_seed(seed)
            self.torch = torch
        else:
            self.g = np.random.default_rng(seed)

    def normal(self, shape, std: float) -> np.ndarray:
        if self.fast:
            t = self.torch.empty(shape, dtype=self.torch.float32)
            t.normal_(0.0, std, generator=self.g)
            return t.numpy()
        return (self.g.standard_normal(shape, dtype=np.float32) * np.float32(std)).astype(np.float32)


def dit_tensors(d: DitDims, seed: int = 0, fast: bool = False) -> Dict[str, np.ndarray]:
    """name -> float32 array in PyTorch shapes."""
    r = _Rng(seed, fast)
    H, I = d.hidden_size, d.intermediate_size
    nq, nkv, hd = d.num_attention_heads, d.num_key_value_heads, d.head_dim
    P, C, A = d.patch_size, d.in_channels, d.audio_acoustic_hidden_dim
    T: Dict[str, np.ndarray] = {}

    def lin(name, out_f, in_f, bias=False, std=None):
        T[name + ".weight"] = r.normal((out_f, in_f), std if std is not None else in_f ** -0.5)
        if bias:
            T[name + ".bias"] = r.normal((out_f,), 0.02)

    def norm(name, n):
        T[name] = (1.0 + r.normal((n,), 0.05)).astype(np.float32)

    T["decoder.proj_in.1.weight"] = r.normal((H, C, P), (C * P) ** -0.5)
    T["decoder.proj_in.1.bias"] = r.normal((H,), 0.02)
    T["decoder.proj_out.1.weight"] = r.normal((H, A, P), H ** -0.5)
    T["decoder.proj_out.1.bias"] = r.normal((A,), 0.02)
    lin("decoder.condition_embedder", H, H, bias=True)
    norm("decoder.norm_out.weight", H)
    T["decoder.scale_shift_table"] = r.normal((1, 2, H), 0.02)
    for te in ("decoder.time_embed", "decoder.time_embed_r"):
        lin(te + ".linear_1", H, 256, bias=True)
        lin(te + ".linear_2", H, H, bias=True)
        # keep the AdaLN modulation O(0.1) so (1+scale) stays well away from 0
        lin(te + ".time_proj", 6 * H, H, bias=True, std=0.1 * H ** -0.5)
    for i in range(d.num_hidden_layers):
        p = f"decoder.layers.{i}."
        norm(p + "self_attn_norm.weight", H)
        norm(p + "cross_attn_norm.weight", H)
        norm(p + "mlp_norm.weight", H)
        for a in ("self_attn", "cross_attn"):
            lin(p + a + ".q_proj", nq * hd, H)
            lin(p + a + ".k_proj", nkv * hd, H)
            lin(p + a + ".v_proj", nkv * hd, H)
            lin(p + a + ".o_proj", H, nq * hd)
            norm(p + a + ".q_norm.weight", hd)
            norm(p + a + ".k_norm.weight", hd)
        lin(p + "mlp.gate_proj", I, H)
        lin(p + "mlp.up_proj", I, H)
        lin(p + "mlp.down_proj", H, I)
        # gate entries sit near 1 so that every branch contributes to the residual stream
        tab = r.normal((1, 6, H), 0.02)
        tab[0, 2] += 1.0
        tab[0, 5] += 1.0
        T[p + "scale_shift_table"] = tab

    def enc_layers(prefix, n):
        for i in range(n):
            q = f"{prefix}layers.{i}."
            norm(q + "input_layernorm.weight", H)
            norm(q + "post_attention_layernorm.weight", H)
            lin(q + "self_attn.q_proj", nq * hd, H)
            lin(q + "self_attn.k_proj", nkv * hd, H)
            lin(q + "self_attn.v_proj", nkv * hd, H)
            lin(q + "self_attn.o_proj", H, nq * hd)
            norm(q + "self_attn.q_norm.weight", hd)
            norm(q + "self_attn.k_norm.weight", hd)
            lin(q + "mlp.gate_proj", I, H)
            lin(q + "mlp.up_proj", I, H)
            lin(q + "mlp.down_proj", H, I)

    if d.num_lyric_encoder_hidden_layers > 0 or d.num_timbre_encoder_hidden_layers > 0:
        lin("encoder.text_projector", H, d.text_hidden_dim)
    if d.num_lyric_encoder_hidden_layers > 0:
        lin("encoder.lyric_encoder.embed_tokens", H, d.text_hidden_dim, bias=True)
        norm("encoder.lyric_encoder.norm.weight", H)
        enc_layers("encoder.lyric_encoder.", d.num_lyric_encoder_hidden_layers)
    if d.num_timbre_encoder_hidden_layers > 0:
        lin("encoder.timbre_encoder.embed_tokens", H, d.timbre_hidden_dim, bias=True)
        norm("encoder.timbre_encoder.norm.weight", H)
        enc_layers("encoder.timbre_encoder.", d.num_timbre_encoder_hidden_layers)
    return T


def qwen_tensors(d: QwenDims, seed: int = 0, fast: bool = False) -> Dict[str, np.ndarray]:
    """Qwen3 tensors under the names the reference loader reads (qwen_model.cpp:425-470)."""
    r = _Rng(seed, fast)
    H, I, hd = d.hidden_size, d.intermediate_size, d.head_dim
    nq, nkv = d.num_attention_heads, d.num_key_value_heads
    T: Dict[str, np.ndarray] = {}
    T["embed_tokens.weight"] = r.normal((d.vocab_size, H), 1.0)
    T["norm.weight"] = (1.0 + r.normal((H,), 0.05)).astype(np.float32)
    for i in range(d.num_hidden_layers):
        p = f"layers.{i}."
        for nm, n in (("input_layernorm.weight", H), ("post_attention_layernorm.weight", H),
                      ("self_attn.q_norm.weight", hd), ("self_attn.k_norm.weight", hd)):
            T[p + nm] = (1.0 + r.normal((n,), 0.05)).astype(np.float32)
        T[p + "self_attn.q_proj.weight"] = r.normal((nq * hd, H), H ** -0.5)
        T[p + "self_attn.k_proj.weight"] = r.normal((nkv * hd, H), H ** -0.5)
        T[p + "self_attn.v_proj.weight"] = r.normal((nkv * hd, H), H ** -0.5)
        T[p + "self_attn.o_proj.weight"] = r.normal((H, nq * hd), (nq * hd) ** -0.5)
        T[p + "mlp.gate_proj.weight"] = r.normal((I, H), H ** -0.5)
        T[p + "mlp.up_proj.weight"] = r.normal((I, H), H ** -0.5)
        T[p + "mlp.down_proj.weight"] = r.normal((H, I), I ** -0.5)
    return T


def make_qwen_fixture(out_dir: str, dims: QwenDims = QWEN_TOY, seed: int = 0, *, safetensors: bool = True,
                      gguf_quant: Optional[str] = None, ggml_lib: Optional[str] = None, fast: bool = False) -> str:
    os.makedirs(out_dir, exist_ok=True)
    with open(os.path.join(out_dir, "config.json"), "w") as f:
        json.dump(dims.config_json(), f, indent=1)
    T = qwen_tensors(dims, seed, fast)
    if safetensors:
        save_safetensors(os.path.join(out_dir, "model.safetensors"), T)
    if gguf_quant:
        save_gguf(os.path.join(out_dir, "model.gguf"), T, gguf_quant, ggml_lib, arch="qwen3", fast=fast)
    return out_dir


# --------------------------------------------------------------------------- GGUF export

class _ChunkQuantizer:
    """ggml_quantize_chunk through ctypes, as the reference exporter does
    (export_safetensors_to_gguf.py:120-151)."""

    def __init__(self, lib_path: str):
        self.lib = ctypes.CDLL(lib_path)
        self.lib.ggml_quantize_chunk.argtypes = [
            ctypes.c_int, ctypes.POINTER(ctypes.c_float), ctypes.c_void_p,
            ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.POINTER(ctypes.c_float)]
        self.lib.ggml_quantize_chunk.restype = ctypes.c_size_t

    def quantize(self, f32_2d: np.ndarray, qtype_value: int, q_byte_row: int) -> np.ndarray:
        f32_2d = np.ascontiguousarray(f32_2d, dtype=np.float32)
        rows, row = f32_2d.shape
        out = np.empty((rows, q_byte_row), dtype=np.uint8)
        n = self.lib.ggml_quantize_chunk(
            qtype_value, f32_2d.ctypes.data_as(ctypes.POINTER(ctypes.c_float)),
            out.ctypes.data_as(ctypes.c_void_p), 0, rows, row, None)
        if int(n) != rows * q_byte_row:
            raise RuntimeError("ggml_quantize_chunk size mismatch")
        return out

    def quantize_parallel(self, f32_2d: np.ndarray, qtype_value: int, q_byte_row: int) -> np.ndarray:
        from concurrent.futures import ThreadPoolExecutor
        rows = f32_2d.shape[0]
        n = min(os.cpu_count() or 1, max(1, rows // 64))
        if n <= 1:
            return self.quantize(f32_2d, qtype_value, q_byte_row)
        cuts = [rows * i // n for i in range(n + 1)]
        with ThreadPoolExecutor(n) as ex:
            parts = list(ex.map(lambda i: self.quantize(f32_2d[cuts[i]:cuts[i + 1]], qtype_value, q_byte_row), range(n)))
        return np.concatenate(parts, axis=0)


def save_gguf(path: str, tensors: Dict[str, np.ndarray], quant: str,
              ggml_lib: Optional[str] = None, arch: str = "acestep", fast: bool = False) -> Dict[str, str]:
    """Write a GGUF twin of ``tensors`` under the exporter's rule. Returns name->type.
    ``fast`` (the 1.6 G-parameter fixtures): every quantised type goes through ggml_quantize_chunk, rows split over
    a thread pool (ctypes drops the GIL); otherwise Q8_0 uses gguf-py's NumPy quantiser."""
    import gguf
    from gguf import quants as gq

    quant = {"Q8": "Q8_0", "Q6": "Q6_K", "Q4": "Q4_K"}.get(quant.upper(), quant.upper())
    qtype = getattr(gguf.GGMLQuantizationType, quant)
    is_q = quant != "F16"
    blk = gguf.GGML_QUANT_SIZES[qtype][0] if is_q else 1
    chunk = _ChunkQuantizer(ggml_lib) if (is_q and ggml_lib) else None

    w = gguf.GGUFWriter(path, arch=arch)
    w.add_type("model")
    w.add_quantization_version(gguf.GGML_QUANT_VERSION)
    kinds: Dict[str, str] = {}
    for name, a in tensors.items():
        a = np.asarray(a, dtype=np.float32)
        if is_q and a.ndim >= 2 and a.shape[-1] % blk == 0:
            rows = int(np.prod(a.shape[:-1]))
            a2 = a.reshape(rows, a.shape[-1])
            brow = gq.quant_shape_to_byte_shape((rows, a.shape[-1]), qtype)[-1]
            if fast and chunk is not None:
                q = chunk.quantize_parallel(a2, int(qtype), int(brow))
            else:
                try:
                    q = gq.quantize(a2, qtype)
                except NotImplementedError:
                    if chunk is None:
                        raise RuntimeError(f"{quant} export needs ggml_lib (ggml_quantize_chunk)")
                    q = chunk.quantize(a2, int(qtype), int(brow))
            q = q.reshape(tuple(a.shape[:-1]) + (q.shape[-1],))
            w.add_tensor(name, q, raw_dtype=qtype)
            kinds[name] = quant
        else:
            w.add_tensor(name, a.astype(np.float16))
            kinds[name] = "F16"
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file(progress=False)
    w.close()
    return kinds


def make_dit_fixture(out_dir: str, dims: DitDims = TOY, seed: int = 0, *,
                     safetensors: bool = True, gguf_quant: Optional[str] = None,
                     ggml_lib: Optional[str] = None, fast: bool = False) -> str:
    """Write a DiT model directory. With ``gguf_quant`` a ``model.gguf`` is written
    (which both loaders prefer over ``model.safetensors``, acestep_dit_model.cpp:47-70)."""
    os.makedirs(out_dir, exist_ok=True)
    with open(os.path.join(out_dir, "config.json"), "w") as f:
        json.dump(dims.config_json(), f, indent=1)
    T = dit_tensors(dims, seed, fast)
    if safetensors:
        save_safetensors(os.path.join(out_dir, "model.safetensors"), T)
    if gguf_quant:
        save_gguf(os.path.join(out_dir, "model.gguf"), T, gguf_quant, ggml_lib, fast=fast)
    return out_dir


# --------------------------------------------------------------------------- VAE weights

def vae_decoder_channels(v: VaeDims) -> List[int]:
    """[C_in of block0, ..., C_out of last block]; Oobleck prepends a 1 to the multiples."""
    m = [1] + list(v.channel_multiples)
    return [v.decoder_channels * x for x in reversed(m)]


def vae_tensors(v: VaeDims, seed: int = 0, fast: bool = False) -> Dict[str, np.ndarray]:
    r = _Rng(seed, fast)
    T: Dict[str, np.ndarray] = {}

    def wn_conv(prefix, d0, d1, k, bias_n, gain=1.0):
        # weight-norm pair; after folding every d0-row has L2 norm g. g is chosen so a unit-RMS
        # input keeps roughly unit-RMS output (variance-preserving): F16 im2col must not overflow.
        vv = r.normal((d0, d1, k), 1.0)
        T[prefix + ".weight_v"] = vv
        T[prefix + ".weight_g"] = (gain * (1.0 + r.normal((d0, 1, 1), 0.05))).astype(np.float32)
        if bias_n:
            T[prefix + ".bias"] = r.normal((bias_n,), 0.02)

    def snake(prefix, c):
        T[prefix + ".alpha"] = r.normal((1, c, 1), 0.1)
        T[prefix + ".beta"] = r.normal((1, c, 1), 0.1)

    def res_unit(prefix, c):
        snake(prefix + ".snake1", c)
        wn_conv(prefix + ".conv1", c, c, 7, c, gain=0.6)
        snake(prefix + ".snake2", c)
        wn_conv(prefix + ".conv2", c, c, 1, c, gain=0.4)

    ch = vae_decoder_channels(v)
    ups = list(reversed(v.downsampling_ratios))
    wn_conv("decoder.conv1", ch[0], v.decoder_input_channels, 7, ch[0], gain=0.9)
    for i, s in enumerate(ups):
        p = f"decoder.block.{i}"
        cin, cout = ch[i], ch[i + 1]
        snake(p + ".snake1", cin)
        # ConvTranspose1d weight is [in, out, k]; weight-norm rows run over d0 = in
        # (acestep_vae_model.cpp:552-565). Each output sample sums 2 taps x cin inputs.
        wn_conv(p + ".conv_t1", cin, cout, 2 * s, cout, gain=0.7 * (cout * 2 * s / (2 * cin)) ** 0.5)
        for j in (1, 2, 3):
            res_unit(f"{p}.res_unit{j}", cout)
    snake("decoder.snake1", ch[-1])
    wn_conv("decoder.conv2", v.audio_channels, ch[-1], 7, 0, gain=0.5)

    # encoder tensors: required to exist by the reference loader (acestep_vae_model.cpp:913-929)
    em = [1] + list(v.channel_multiples)
    ech = [v.encoder_hidden_size * x for x in em]
    wn_conv("encoder.conv1", ech[0], v.audio_channels, 7, ech[0], gain=0.9)
    for i, s in enumerate(v.downsampling_ratios):
        p = f"encoder.block.{i}"
        cin, cout = ech[i], ech[i + 1]
        for j in (1, 2, 3):
            res_unit(f"{p}.res_unit{j}", cin)
        snake(p + ".snake1", cin)
        wn_conv(p + ".conv1", cout, cin, 2 * s, cout, gain=0.7)
    snake("encoder.snake1", ech[-1])
    wn_conv("encoder.conv2", v.encoder_hidden_size, ech[-1], 3, v.encoder_hidden_size, gain=0.5)
    return T


def make_vae_fixture(out_dir: str, dims: VaeDims = VAE_TOY, seed: int = 0, fast: bool = False) -> str:
    os.makedirs(out_dir, exist_ok=True)
    with open(os.path.join(out_dir, "config.json"), "w") as f:
        json.dump(dims.config_json(), f, indent=1)
    save_safetensors(os.path.join(out_dir, "diffusion_pytorch_model.safetensors"),
                     vae_tensors(dims, seed, fast))
    return out_dir


# --------------------------------------------------------------------------- synthetic inputs

def dit_inputs(dims: DitDims, seq_len: int, enc_len: int, seed: int = 1234) -> Dict[str, np.ndarray]:
    """Seeded inputs in the shape compare_dit.py:576-579 uses; context = [N(0,1) x 64 | 1.0 x 64]
    (silence-latent half + chunk-mask half, acestep_ggml.cpp:1946-2040)."""
    g = np.random.default_rng(seed)
    A = dims.audio_acoustic_hidden_dim
    ctx_dim = dims.in_channels - A
    hidden = g.standard_normal((seq_len, A), dtype=np.float32)
    ctx = np.ones((seq_len, ctx_dim), dtype=np.float32)
    ctx[:, : ctx_dim // 2] = g.standard_normal((seq_len, ctx_dim // 2), dtype=np.float32)
    enc = g.standard_normal((max(enc_len, 1), dims.hidden_size), dtype=np.float32)[:enc_len]
    return {"hidden_states": hidden, "context_latents": ctx, "encoder_hidden_states": enc}


def vae_latents(dims: VaeDims, n_frames: int, seed: int = 7) -> np.ndarray:
    g = np.random.default_rng(seed)
    return g.standard_normal((n_frames, dims.decoder_input_channels), dtype=np.float32)
