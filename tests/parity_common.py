"""Shared helpers for the parity tests: seeded fixtures on disk, reference (oracle/_ref) runs, distance metrics.
Metrics follow the reference's own compare tools (tools/compare_dit.py:576-627: MAE / max-abs / cosine;
tools/eval_quant_prompt_pipeline.py:197-225: SNR)."""
import os

import numpy as np

from tools import fixtures as fx
from oracle import refbridge


def metrics(a: np.ndarray, ref: np.ndarray) -> dict:
    a = np.asarray(a, dtype=np.float64).ravel()
    r = np.asarray(ref, dtype=np.float64).ravel()
    d = a - r
    return {
        "rel_l2": float(np.linalg.norm(d) / (np.linalg.norm(r) + 1e-30)),
        "cos": float(a @ r / (np.linalg.norm(a) * np.linalg.norm(r) + 1e-30)),
        "mae": float(np.abs(d).mean()),
        "max_abs": float(np.abs(d).max()),
        "snr_db": float(10 * np.log10((r @ r + 1e-30) / (d @ d + 1e-30))),
    }


def make_dit_dir(root: str, dims=fx.TOY, seed: int = 0, quant: str = None) -> str:
    """safetensors fixture, or (quant in F16/Q8_0/Q4_K/Q6_K) a GGUF-only directory."""
    name = f"dit_{dims.hidden_size}_{dims.num_hidden_layers}_{quant or 'st'}"
    d = os.path.join(root, name)
    if os.path.exists(os.path.join(d, "config.json")):
        return d
    if quant:
        fx.make_dit_fixture(d, dims, seed, safetensors=False, gguf_quant=quant,
                            ggml_lib=refbridge.ref_lib() if refbridge.ref_available() else None)
    else:
        fx.make_dit_fixture(d, dims, seed)
    return d


def make_vae_dir(root: str, dims=fx.VAE_TOY, seed: int = 0) -> str:
    d = os.path.join(root, f"vae_{dims.decoder_channels}")
    if not os.path.exists(os.path.join(d, "config.json")):
        fx.make_vae_fixture(d, dims, seed)
    return d


def ref_dit_forward(dit_dir, inputs, t, t_r, max_layers=None, threads=4, buffer_bytes=1 << 30, masks=None):
    env = {}
    if max_layers:
        env["ACE_GGML_DIT_MAX_LAYERS"] = str(max_layers)
    call = {"op": "dit_forward", "t": t, "t_r": t_r, "out": "v"}
    inp = dict(inputs)
    if masks:
        if masks.get("mask") is not None:
            inp["attention_mask"] = masks["mask"]
            call["mask"] = "attention_mask"
        if masks.get("enc_mask") is not None:
            inp["encoder_attention_mask"] = masks["enc_mask"]
            call["enc_mask"] = "encoder_attention_mask"
    outs, tim = refbridge.run_ref([call], inp, dit_dir=dit_dir, threads=threads, compute_buffer_bytes=buffer_bytes, env=env)
    return outs["v"], tim


def dequantised_gguf_tensors(path: str) -> dict:
    """name -> float32 array (PyTorch shapes) of a GGUF file, quantised tensors dequantised with gguf.quants.dequantize
    (the numpy port of ggml-quants.c).  Used as the "truth for the same quantised weights" calibrator."""
    import gguf
    from gguf import quants as gq
    out = {}
    for t in gguf.GGUFReader(path).tensors:
        shape = [int(x) for x in reversed(t.shape)]
        if t.tensor_type in (gguf.GGMLQuantizationType.F32, gguf.GGMLQuantizationType.F16):
            out[t.name] = np.asarray(t.data, dtype=np.float32).reshape(shape)
        else:
            out[t.name] = gq.dequantize(np.asarray(t.data), t.tensor_type).astype(np.float32).reshape(shape)
    return out


# ---------------------------------------------------------------- real-dimension fixtures (BASELINE configs)

def fixtures_root() -> str:
    r = os.environ.get("ACE_B200_FIXTURES", "/tmp/ace_b200_fixtures")
    os.makedirs(r, exist_ok=True)
    return r


def turbo_dit_dir(quant: str) -> str:
    """turbo-dimension DiT as a GGUF-only directory (seed 0, torch CPU generator; shared with bench.py)."""
    d = os.path.join(fixtures_root(), f"turbo_{quant}")
    if not os.path.exists(os.path.join(d, "model.gguf")):
        tmp = d + f".tmp{os.getpid()}"
        lib = refbridge.ref_lib() if (quant != "F16" and refbridge.ref_available()) else None
        if quant != "F16" and lib is None:
            raise RuntimeError("quantised turbo fixtures need ggml_quantize_chunk from oracle/_ref")
        fx.make_dit_fixture(tmp, fx.TURBO, 0, safetensors=False, gguf_quant=quant, ggml_lib=lib, fast=True)
        try:
            os.rename(tmp, d)
        except OSError:
            pass
    return d


def full_vae_dir() -> str:
    d = os.path.join(fixtures_root(), "vae_full")
    if not os.path.exists(os.path.join(d, "config.json")):
        fx.make_vae_fixture(d, fx.VAE_FULL, 0, fast=True)
    return d


def fixture_checksums(dit_dir: str) -> dict:
    """crc32 of the raw bytes of three tensors of model.gguf: tells 'the regenerated fixture differs from the one the
    golden was made on' apart from 'the kernel differs'."""
    import zlib

    import gguf
    want = ("decoder.layers.0.self_attn.q_proj.weight", "decoder.time_embed.time_proj.weight", "decoder.proj_in.1.weight")
    out = {}
    for t in gguf.GGUFReader(os.path.join(dit_dir, "model.gguf")).tensors:
        if t.name in want:
            out["wsum_" + t.name.replace(".", "_")] = np.int64(zlib.crc32(np.ascontiguousarray(t.data).tobytes()))
    return out


def shift_schedule(shift: float, n: int) -> np.ndarray:
    u = 1.0 - np.arange(n, dtype=np.float64) / n
    return (shift * u / (1.0 + (shift - 1.0) * u)).astype(np.float32)
