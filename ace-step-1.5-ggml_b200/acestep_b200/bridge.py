"""ctypes bridge to ``libacestep_b200.so`` — the host-side mirror of the reference's plugin seam.

The acestep pipeline reaches the ggml library through ``GGMLCAPIBridge`` and two monkey-patch installers
(reference: ``scripts/run_non_ggml_real_case.py:135-354`` bridge, ``:445-538`` DiT hook, ``:541-661`` VAE hook).
``B200Bridge`` keeps that class's method names, argument meaning and error behaviour (``RuntimeError`` carrying
``ace_ggml_last_error``) so the same hooks work when handed this object, and adds the batched / device-resident
entry points of ``include/acestep_b200.h``.

There is no CPU fallback: if the shared library is missing, or no sm_100 device is present, construction raises.
"""
from __future__ import annotations

import ctypes
import os
from pathlib import Path
from typing import Optional, Sequence

import numpy as np

_HERE = Path(__file__).resolve().parent
DEFAULT_LIB = _HERE / "lib" / "libacestep_b200.so"

_F = ctypes.POINTER(ctypes.c_float)
_I = ctypes.POINTER(ctypes.c_int32)
_P = ctypes.c_void_p

STATUS_NAMES = {0: "OK", 1: "ERR", 2: "ERR_INVALID_ARG", 3: "ERR_IO", 4: "ERR_UNSUPPORTED"}


class AceInitParams(ctypes.Structure):
    """``ace_ggml_init_params`` (acestep_ggml.h:31-35)."""
    _fields_ = [("n_threads", ctypes.c_int32), ("use_metal", ctypes.c_int32),
                ("compute_buffer_bytes", ctypes.c_size_t)]


# every symbol include/acestep_ggml.h and include/acestep_b200.h declare: (restype, argtypes)
SIGNATURES = {
    "ace_ggml_create": (ctypes.c_int, [ctypes.POINTER(AceInitParams), ctypes.POINTER(_P)]),
    "ace_ggml_destroy": (None, [_P]),
    "ace_ggml_last_error": (ctypes.c_char_p, [_P]),
    "ace_ggml_load_lm": (ctypes.c_int, [_P, ctypes.c_char_p]),
    "ace_ggml_load_text_encoder": (ctypes.c_int, [_P, ctypes.c_char_p]),
    "ace_ggml_load_dit": (ctypes.c_int, [_P, ctypes.c_char_p]),
    "ace_ggml_load_vae": (ctypes.c_int, [_P, ctypes.c_char_p]),
    "ace_ggml_vae_get_info": (ctypes.c_int, [_P, _I, _I, _I]),
    "ace_ggml_vae_decode": (ctypes.c_int, [_P, _F, ctypes.c_int32, _F, ctypes.c_size_t]),
    "ace_ggml_vae_encode": (ctypes.c_int, [_P, _F, ctypes.c_int32, _F, ctypes.c_size_t]),
    "ace_ggml_text_encoder_forward": (ctypes.c_int, [_P, _I, ctypes.c_int32, _F, ctypes.c_size_t]),
    "ace_ggml_text_encoder_forward_masked": (ctypes.c_int, [_P, _I, _I, ctypes.c_int32, _F, ctypes.c_size_t]),
    "ace_ggml_text_encoder_forward_embeddings": (ctypes.c_int, [_P, _I, ctypes.c_int32, _F, ctypes.c_size_t]),
    "ace_ggml_text_encoder_forward_layers": (ctypes.c_int, [_P, _I, _I, ctypes.c_int32, ctypes.c_int32,
                                                            ctypes.c_int32, _F, ctypes.c_size_t]),
    "ace_ggml_dit_forward": (ctypes.c_int, [_P, _F, _F, _F, _I, _I, ctypes.c_int32, ctypes.c_int32,
                                            ctypes.c_float, ctypes.c_float, _F, ctypes.c_size_t]),
    "ace_ggml_generate_audio_simple": (ctypes.c_int, [_P, _I, ctypes.c_int32, ctypes.c_int32, ctypes.c_float,
                                                      ctypes.c_int32, _F, ctypes.c_size_t, _I, _I]),
    "ace_ggml_generate_audio_style_lyric_simple": (ctypes.c_int, [_P, _I, ctypes.c_int32, _I, ctypes.c_int32,
                                                                  ctypes.c_int32, ctypes.c_float, ctypes.c_int32,
                                                                  _F, ctypes.c_size_t, _I, _I]),
    "ace_ggml_generate_audio_style_lyric_timbre_simple": (ctypes.c_int, [_P, _I, ctypes.c_int32, _I, ctypes.c_int32,
                                                                         _F, _I, ctypes.c_int32, ctypes.c_int32,
                                                                         ctypes.c_int32, ctypes.c_float,
                                                                         ctypes.c_int32, _F, ctypes.c_size_t, _I, _I]),
    "ace_b200_dit_forward_batched": (ctypes.c_int, [_P, _F, _F, _F, _I, _I, ctypes.c_int32, ctypes.c_int32,
                                                    ctypes.c_int32, _F, _F, _F, ctypes.c_size_t]),
    "ace_b200_generate_latents": (ctypes.c_int, [_P, _F, _F, _F, _F, _I, _I, ctypes.c_int32, ctypes.c_int32,
                                                 ctypes.c_int32, _F, ctypes.c_int32, ctypes.c_float, _F,
                                                 ctypes.c_size_t]),
    "ace_b200_vae_decode_batched": (ctypes.c_int, [_P, _F, ctypes.c_int32, ctypes.c_int32, _F, ctypes.c_size_t]),
    "ace_b200_generate_audio": (ctypes.c_int, [_P, _F, _F, _F, _F, _I, _I, ctypes.c_int32, ctypes.c_int32,
                                               ctypes.c_int32, _F, ctypes.c_int32, ctypes.c_float, _F, _F,
                                               ctypes.c_size_t]),
    "ace_b200_generate_audio_from_encoder": (ctypes.c_int, [_P, _F, _I, ctypes.c_int32, ctypes.c_int32, ctypes.c_float,
                                                            ctypes.c_int32, _F, ctypes.c_size_t, _I, _I]),
    "ace_b200_lyric_encoder_forward": (ctypes.c_int, [_P, _F, ctypes.c_int32, _F, ctypes.c_size_t]),
    "ace_b200_timbre_encoder_forward": (ctypes.c_int, [_P, _F, ctypes.c_int32, _F, ctypes.c_size_t]),
    "ace_b200_create_on_device": (ctypes.c_int, [ctypes.POINTER(AceInitParams), ctypes.c_int32, ctypes.POINTER(_P)]),
    "ace_b200_peer_buffer_create": (ctypes.c_int, [_P, ctypes.c_size_t, _P]),
    "ace_b200_peer_buffer_open": (ctypes.c_int, [_P, _P, ctypes.c_size_t]),
    "ace_b200_peer_buffer_attach": (ctypes.c_int, [_P, _P]),
    "ace_b200_vae_decode_window": (ctypes.c_int, [_P, _F, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_int64]),
    "ace_b200_peer_buffer_read": (ctypes.c_int, [_P, _F, ctypes.c_size_t]),
    "ace_b200_shift_schedule": (ctypes.c_int, [ctypes.c_float, ctypes.c_int32, _F]),
    "ace_b200_last_device_ms": (ctypes.c_double, [_P]),
    "ace_b200_kernel_launches": (ctypes.c_uint64, []),
    "ace_b200_model_info": (ctypes.c_char_p, [_P]),
    "ace_b200_profile_enable": (None, [ctypes.c_int32]),
    "ace_b200_profile_collect": (None, [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double),
                                        ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int64)]),
    "ace_b200_debug_gemm": (ctypes.c_int, [_P, _F, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, _P, ctypes.c_int32,
                                           ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                           ctypes.c_int32, _F, _F, ctypes.c_int32, ctypes.c_int32, _F,
                                           ctypes.c_size_t]),
    "ace_b200_debug_attention": (ctypes.c_int, [_P, _F, _F, _F, _I, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                                ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, _F, ctypes.c_size_t]),
    "ace_b200_attention_noalloc": (ctypes.c_int, [_P, _F, _F, _F, _I, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                                  ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, _F, ctypes.c_size_t]),
    "ace_b200_debug_quantize_rows": (ctypes.c_int, [_F, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, _P, ctypes.c_size_t]),
    "ace_b200_debug_probe_tensor_file": (ctypes.c_int64, [ctypes.c_char_p, ctypes.c_int32, ctypes.c_char_p, ctypes.c_size_t]),
    "ace_b200_debug_pack_roundtrip": (ctypes.c_int, [_P, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, _F, _F]),
}


def load_library(lib_path: Optional[os.PathLike] = None) -> ctypes.CDLL:
    """dlopen the library and bind every declared symbol (raises if the library or a symbol is missing)."""
    path = Path(lib_path) if lib_path else Path(os.environ.get("ACE_B200_LIB", DEFAULT_LIB))
    if not path.exists():
        raise RuntimeError(f"{path} not found: build it with `make -C ace-step-1.5-ggml_b200/csrc` "
                           "(there is no CPU fallback)")
    lib = ctypes.CDLL(str(path))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the export is missing
        fn.restype = res
        fn.argtypes = args
    return lib


def _f(a: Optional[np.ndarray]):
    return a.ctypes.data_as(_F) if a is not None else None


def _i(a: Optional[np.ndarray]):
    return a.ctypes.data_as(_I) if a is not None else None


def _f32(a, ndim=None) -> Optional[np.ndarray]:
    if a is None:
        return None
    a = np.ascontiguousarray(a, dtype=np.float32)
    if ndim is not None and a.ndim != ndim:
        raise ValueError(f"expected a {ndim}-D array, got shape {a.shape}")
    return a


def _i32(a) -> Optional[np.ndarray]:
    return None if a is None else np.ascontiguousarray(a, dtype=np.int32)


class B200Bridge:
    """Same surface as the reference ``GGMLCAPIBridge`` (run_non_ggml_real_case.py:135-354)."""

    def __init__(self, lib_path: Optional[os.PathLike] = None, n_threads: int = 1, compute_buffer_mb: int = 0,
                 use_metal: bool = False, device: Optional[int] = None):
        """device=None: ACE_B200_DEVICE, else the current CUDA device (ace_ggml_create); an int names the GPU explicitly
        (ace_b200_create_on_device) — several bridges on several GPUs can live in one process."""
        self.lib = load_library(lib_path)
        self.ctx = _P()
        prm = AceInitParams(int(n_threads), 1 if use_metal else 0, int(compute_buffer_mb) * 1024 * 1024)
        if device is None:
            st = self.lib.ace_ggml_create(ctypes.byref(prm), ctypes.byref(self.ctx))
        else:
            st = self.lib.ace_b200_create_on_device(ctypes.byref(prm), int(device), ctypes.byref(self.ctx))
        if st != 0 or not self.ctx:
            self.ctx = _P()
            raise RuntimeError(f"ace_ggml_create failed with status {st} ({STATUS_NAMES.get(st, '?')}): "
                               "an sm_100 CUDA device is required")
        self.latent_channels = 0
        self.audio_channels = 0
        self.hop_length = 0
        self.text_hidden = 0
        self.dit_dims = None

    # ------------------------------------------------------------------ plumbing
    def _last_error(self) -> str:
        msg = self.lib.ace_ggml_last_error(self.ctx)
        return msg.decode("utf-8", "replace") if msg else ""

    def _ensure_ok(self, status: int, where: str) -> None:
        if status != 0:
            raise RuntimeError(f"{where} failed: status={status} ({STATUS_NAMES.get(status, '?')}) {self._last_error()}")

    def close(self) -> None:
        if getattr(self, "ctx", None):
            self.lib.ace_ggml_destroy(self.ctx)
            self.ctx = _P()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # ------------------------------------------------------------------ loads
    def load_text_encoder(self, model_dir) -> None:
        self._ensure_ok(self.lib.ace_ggml_load_text_encoder(self.ctx, str(model_dir).encode()), "ace_ggml_load_text_encoder")
        import json
        root = Path(model_dir).parent if str(model_dir).endswith(".gguf") else Path(model_dir)
        self.text_hidden = int(json.load(open(root / "config.json"))["hidden_size"])

    def load_dit(self, model_dir) -> None:
        self._ensure_ok(self.lib.ace_ggml_load_dit(self.ctx, str(model_dir).encode()), "ace_ggml_load_dit")

    def load_vae(self, model_dir) -> None:
        self._ensure_ok(self.lib.ace_ggml_load_vae(self.ctx, str(model_dir).encode()), "ace_ggml_load_vae")
        a, b, c = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int32()
        self._ensure_ok(self.lib.ace_ggml_vae_get_info(self.ctx, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c)),
                        "ace_ggml_vae_get_info")
        self.latent_channels, self.audio_channels, self.hop_length = a.value, b.value, c.value

    # ------------------------------------------------------------------ text encoder / VAE encode / self-contained sampler
    def text_encoder_forward(self, token_ids, attention_mask=None, n_layers: Optional[int] = None, apply_final_norm: bool = True,
                             hidden_size: Optional[int] = None) -> np.ndarray:
        """token ids [n] -> hidden states [n, hidden] (causal; optional key mask / layer truncation).  hidden_size defaults to the
        value remembered by load_text_encoder."""
        ids = _i32(token_ids)
        n = int(ids.shape[0])
        H = int(hidden_size or self.text_hidden)
        out = np.empty((n, H), dtype=np.float32)
        am = _i32(attention_mask)
        if n_layers is not None:
            st = self.lib.ace_ggml_text_encoder_forward_layers(self.ctx, _i(ids), _i(am), n, int(n_layers), 1 if apply_final_norm else 0,
                                                               _f(out), out.nbytes)
            where = "ace_ggml_text_encoder_forward_layers"
        elif am is not None:
            st = self.lib.ace_ggml_text_encoder_forward_masked(self.ctx, _i(ids), _i(am), n, _f(out), out.nbytes)
            where = "ace_ggml_text_encoder_forward_masked"
        else:
            st = self.lib.ace_ggml_text_encoder_forward(self.ctx, _i(ids), n, _f(out), out.nbytes)
            where = "ace_ggml_text_encoder_forward"
        self._ensure_ok(st, where)
        return out

    def text_encoder_embeddings(self, token_ids, hidden_size: Optional[int] = None) -> np.ndarray:
        ids = _i32(token_ids)
        n = int(ids.shape[0])
        out = np.empty((n, int(hidden_size or self.text_hidden)), dtype=np.float32)
        self._ensure_ok(self.lib.ace_ggml_text_encoder_forward_embeddings(self.ctx, _i(ids), n, _f(out), out.nbytes),
                        "ace_ggml_text_encoder_forward_embeddings")
        return out

    def vae_encode_tfirst(self, audio_tfirst: np.ndarray) -> np.ndarray:
        """audio [n_samples, channels] -> posterior-mean latents [frames, latent_channels]"""
        a = _f32(audio_tfirst, 2)
        n = int(a.shape[0])
        out = np.empty((max(n // self.hop_length, 1) + 2, self.latent_channels), dtype=np.float32)
        self._ensure_ok(self.lib.ace_ggml_vae_encode(self.ctx, _f(a), n, _f(out), out.nbytes), "ace_ggml_vae_encode")
        return out[: n // self.hop_length]

    def lyric_encoder_forward(self, lyric_embeds: np.ndarray, hidden_size: int) -> np.ndarray:
        x = _f32(lyric_embeds, 2)
        out = np.empty((x.shape[0], hidden_size), np.float32)
        self._ensure_ok(self.lib.ace_b200_lyric_encoder_forward(self.ctx, _f(x), int(x.shape[0]), _f(out), out.nbytes), "ace_b200_lyric_encoder_forward")
        return out

    def timbre_encoder_forward(self, refer_states: np.ndarray, hidden_size: int) -> np.ndarray:
        x = _f32(refer_states, 2)
        out = np.empty((hidden_size,), np.float32)
        self._ensure_ok(self.lib.ace_b200_timbre_encoder_forward(self.ctx, _f(x), int(x.shape[0]), _f(out), out.nbytes), "ace_b200_timbre_encoder_forward")
        return out

    def _audio_out(self, seq_len: int):
        return np.zeros((seq_len * self.hop_length, self.audio_channels), dtype=np.float32), ctypes.c_int32(), ctypes.c_int32()

    def generate_audio_from_encoder(self, enc, enc_mask, seq_len: int, shift: float = 3.0, seed: int = 0) -> np.ndarray:
        e = _f32(enc, 2)
        em = _i32(enc_mask)
        out, ns, nc = self._audio_out(seq_len)
        st = self.lib.ace_b200_generate_audio_from_encoder(self.ctx, _f(e), _i(em), int(e.shape[0]), int(seq_len), float(shift), int(seed),
                                                           _f(out), out.nbytes, ctypes.byref(ns), ctypes.byref(nc))
        self._ensure_ok(st, "ace_b200_generate_audio_from_encoder")
        return out[: ns.value]

    def generate_audio_simple(self, token_ids, seq_len: int, shift: float = 3.0, seed: int = 0) -> np.ndarray:
        ids = _i32(token_ids)
        out, ns, nc = self._audio_out(seq_len)
        st = self.lib.ace_ggml_generate_audio_simple(self.ctx, _i(ids), int(ids.shape[0]), int(seq_len), float(shift), int(seed), _f(out),
                                                     out.nbytes, ctypes.byref(ns), ctypes.byref(nc))
        self._ensure_ok(st, "ace_ggml_generate_audio_simple")
        return out[: ns.value]

    def generate_audio_style_lyric_timbre(self, style_ids, lyric_ids, refer=None, seq_len: int = 0, shift: float = 3.0,
                                          seed: int = 0) -> np.ndarray:
        """refer: [n_refer, refer_len, timbre_dim] or None (then the style+lyric entry point is used)"""
        sid = _i32(style_ids) if style_ids is not None else None
        lid = _i32(lyric_ids) if lyric_ids is not None else None
        ns_, nl_ = (0 if sid is None else int(sid.shape[0])), (0 if lid is None else int(lid.shape[0]))
        out, ns, nc = self._audio_out(seq_len)
        if refer is None:
            st = self.lib.ace_ggml_generate_audio_style_lyric_simple(self.ctx, _i(sid), ns_, _i(lid), nl_, int(seq_len), float(shift), int(seed),
                                                                     _f(out), out.nbytes, ctypes.byref(ns), ctypes.byref(nc))
            where = "ace_ggml_generate_audio_style_lyric_simple"
        else:
            r = _f32(refer, 3)
            order = np.zeros((r.shape[0],), np.int32)
            st = self.lib.ace_ggml_generate_audio_style_lyric_timbre_simple(self.ctx, _i(sid), ns_, _i(lid), nl_, _f(r), _i(order), int(r.shape[0]),
                                                                            int(r.shape[1]), int(seq_len), float(shift), int(seed), _f(out),
                                                                            out.nbytes, ctypes.byref(ns), ctypes.byref(nc))
            where = "ace_ggml_generate_audio_style_lyric_timbre_simple"
        self._ensure_ok(st, where)
        return out[: ns.value]

    # ------------------------------------------------------------------ reference-shaped calls (one sample)
    def vae_decode_tfirst(self, latents_tfirst: np.ndarray) -> np.ndarray:
        """latents [T, C] -> audio [T*hop, channels]."""
        if self.audio_channels <= 0 or self.hop_length <= 0:
            raise RuntimeError("VAE not initialized in bridge")
        lat = _f32(latents_tfirst, 2)
        n = int(lat.shape[0])
        out = np.empty((n * self.hop_length, self.audio_channels), dtype=np.float32)
        self._ensure_ok(self.lib.ace_ggml_vae_decode(self.ctx, _f(lat), n, _f(out), out.nbytes), "ace_ggml_vae_decode")
        return out

    def dit_forward_tfirst(self, hidden_states_tfirst, context_latents_tfirst, encoder_hidden_states_tfirst,
                           attention_mask, encoder_attention_mask, timestep: float, timestep_r: float) -> np.ndarray:
        """hidden [T, D], context [T, Ctx], encoder states [L, H], masks [T] / [L] (or None) -> v [T, D]."""
        hs = _f32(hidden_states_tfirst, 2)
        cx = _f32(context_latents_tfirst, 2)
        enc = _f32(encoder_hidden_states_tfirst, 2) if encoder_hidden_states_tfirst is not None else None
        am, eam = _i32(attention_mask), _i32(encoder_attention_mask)
        T = int(hs.shape[0])
        E = int(enc.shape[0]) if enc is not None else 0
        out = np.empty_like(hs)
        st = self.lib.ace_ggml_dit_forward(self.ctx, _f(hs), _f(cx), _f(enc) if E else None, _i(am), _i(eam), T, E,
                                           float(timestep), float(timestep_r), _f(out), out.nbytes)
        self._ensure_ok(st, "ace_ggml_dit_forward")
        return out

    # ------------------------------------------------------------------ batched / device-resident extensions
    def dit_forward_batched(self, hidden, context, enc, attention_mask, encoder_attention_mask,
                            timesteps: Sequence[float], timesteps_r: Sequence[float],
                            out: Optional[np.ndarray] = None) -> np.ndarray:
        """B samples per call; `out` (optional, e.g. pinned memory) receives the velocities [B, T, D]."""
        hs, cx = _f32(hidden, 3), _f32(context, 3)
        en = _f32(enc, 3) if enc is not None else None
        am, eam = _i32(attention_mask), _i32(encoder_attention_mask)
        B, T = int(hs.shape[0]), int(hs.shape[1])
        E = int(en.shape[1]) if en is not None else 0
        t = _f32(np.broadcast_to(np.asarray(timesteps, dtype=np.float32), (B,)))
        tr = _f32(np.broadcast_to(np.asarray(timesteps_r, dtype=np.float32), (B,)))
        if out is None:
            out = np.empty_like(hs)
        elif out.dtype != np.float32 or out.shape != hs.shape or not out.flags.c_contiguous:
            raise ValueError("out must be a C-contiguous float32 array shaped like hidden")
        st = self.lib.ace_b200_dit_forward_batched(self.ctx, _f(hs), _f(cx), _f(en) if E else None, _i(am), _i(eam),
                                                   B, T, E, _f(t), _f(tr), _f(out), out.nbytes)
        self._ensure_ok(st, "ace_b200_dit_forward_batched")
        return out

    def shift_schedule(self, shift: float, n_steps: int) -> np.ndarray:
        out = np.empty((n_steps,), dtype=np.float32)
        self._ensure_ok(self.lib.ace_b200_shift_schedule(float(shift), int(n_steps), _f(out)), "ace_b200_shift_schedule")
        return out

    def generate_latents(self, noise, context, enc, uncond_enc=None, attention_mask=None, encoder_attention_mask=None,
                         schedule=None, guidance: float = 1.0) -> np.ndarray:
        x, cx = _f32(noise, 3), _f32(context, 3)
        en = _f32(enc, 3) if enc is not None else None
        un = _f32(uncond_enc, 3) if uncond_enc is not None else None
        am, eam = _i32(attention_mask), _i32(encoder_attention_mask)
        sch = _f32(schedule, 1)
        B, T = int(x.shape[0]), int(x.shape[1])
        E = int(en.shape[1]) if en is not None else 0
        out = np.empty_like(x)
        st = self.lib.ace_b200_generate_latents(self.ctx, _f(x), _f(cx), _f(en) if E else None, _f(un), _i(am), _i(eam),
                                                B, T, E, _f(sch), int(sch.shape[0]), float(guidance), _f(out), out.nbytes)
        self._ensure_ok(st, "ace_b200_generate_latents")
        return out

    def vae_decode_time_split(self, latents_tfirst: np.ndarray, rank: int, world: int, halo: int = 16) -> np.ndarray:
        """this rank's slice of a time-split decode of one clip (shard.time_split_plan): decodes the halo'd window and
        returns only the core samples [(c1-c0)*hop, audio_channels]; stitch with shard.gather_time_slices."""
        from . import shard
        lat = _f32(latents_tfirst, 2)
        w0, w1, c0, c1 = shard.time_split_plan(lat.shape[0], rank, world, halo)
        if c1 <= c0:
            return np.zeros((0, self.audio_channels), np.float32)
        audio = self.vae_decode_tfirst(lat[w0:w1])
        s0 = (c0 - w0) * self.hop_length
        return np.ascontiguousarray(audio[s0:s0 + (c1 - c0) * self.hop_length])

    # ---- time-split decode with peer stores (include/acestep_b200.h): the assembled clip lives on the owner's GPU
    def peer_buffer_create(self, n_frames: int) -> bytes:
        """owner: allocate the assembled-clip buffer for n_frames latent frames; returns its 64-byte CUDA IPC handle"""
        h = (ctypes.c_ubyte * 64)()
        nbytes = int(n_frames) * self.hop_length * self.audio_channels * 4
        self._ensure_ok(self.lib.ace_b200_peer_buffer_create(self.ctx, nbytes, ctypes.cast(h, _P)), "ace_b200_peer_buffer_create")
        return bytes(h)

    def peer_buffer_open(self, handle: bytes, n_frames: int) -> None:
        h = (ctypes.c_ubyte * 64).from_buffer_copy(handle)
        nbytes = int(n_frames) * self.hop_length * self.audio_channels * 4
        self._ensure_ok(self.lib.ace_b200_peer_buffer_open(self.ctx, ctypes.cast(h, _P), nbytes), "ace_b200_peer_buffer_open")

    def peer_buffer_attach(self, owner: "B200Bridge") -> None:
        self._ensure_ok(self.lib.ace_b200_peer_buffer_attach(self.ctx, owner.ctx), "ace_b200_peer_buffer_attach")

    def vae_decode_window(self, latents_window: np.ndarray, core_frame0: int, core_frames: int, dst_frame: int) -> None:
        lat = _f32(latents_window, 2)
        self._ensure_ok(self.lib.ace_b200_vae_decode_window(self.ctx, _f(lat), int(lat.shape[0]), int(core_frame0), int(core_frames),
                                                            int(dst_frame)), "ace_b200_vae_decode_window")

    def peer_buffer_read(self, n_frames: int) -> np.ndarray:
        out = np.empty((int(n_frames) * self.hop_length, self.audio_channels), dtype=np.float32)
        self._ensure_ok(self.lib.ace_b200_peer_buffer_read(self.ctx, _f(out), out.nbytes), "ace_b200_peer_buffer_read")
        return out

    def vae_decode_batched(self, latents) -> np.ndarray:
        lat = _f32(latents, 3)
        B, T = int(lat.shape[0]), int(lat.shape[1])
        out = np.empty((B, T * self.hop_length, self.audio_channels), dtype=np.float32)
        self._ensure_ok(self.lib.ace_b200_vae_decode_batched(self.ctx, _f(lat), B, T, _f(out), out.nbytes),
                        "ace_b200_vae_decode_batched")
        return out

    def generate_audio(self, noise, context, enc, uncond_enc=None, attention_mask=None, encoder_attention_mask=None,
                       schedule=None, guidance: float = 1.0, out_audio: Optional[np.ndarray] = None,
                       return_latents: bool = False):
        x, cx = _f32(noise, 3), _f32(context, 3)
        en = _f32(enc, 3) if enc is not None else None
        un = _f32(uncond_enc, 3) if uncond_enc is not None else None
        am, eam = _i32(attention_mask), _i32(encoder_attention_mask)
        sch = _f32(schedule, 1)
        B, T = int(x.shape[0]), int(x.shape[1])
        E = int(en.shape[1]) if en is not None else 0
        audio = out_audio if out_audio is not None else np.empty((B, T * self.hop_length, self.audio_channels), np.float32)
        lat = np.empty_like(x) if return_latents else None
        st = self.lib.ace_b200_generate_audio(self.ctx, _f(x), _f(cx), _f(en) if E else None, _f(un), _i(am), _i(eam),
                                              B, T, E, _f(sch), int(sch.shape[0]), float(guidance), _f(lat), _f(audio),
                                              audio.nbytes)
        self._ensure_ok(st, "ace_b200_generate_audio")
        return (audio, lat) if return_latents else audio

    # ------------------------------------------------------------------ measurement
    def last_device_ms(self) -> float:
        return float(self.lib.ace_b200_last_device_ms(self.ctx))

    def kernel_launches(self) -> int:
        return int(self.lib.ace_b200_kernel_launches())

    def model_info(self) -> str:
        return self.lib.ace_b200_model_info(self.ctx).decode()

    def profile_enable(self, on: bool) -> None:
        self.lib.ace_b200_profile_enable(1 if on else 0)

    def profile_collect(self) -> dict:
        """per-family totals since profile_enable(True): {family: {ms, flops, bytes, launches}}"""
        ms, fl, by = (ctypes.c_double * 4)(), (ctypes.c_double * 4)(), (ctypes.c_double * 4)()
        n = (ctypes.c_int64 * 4)()
        self.lib.ace_b200_profile_collect(ms, fl, by, n)
        names = ("gemm", "attention", "elementwise", "vae_tail")
        return {k: {"ms": ms[i], "flops": fl[i], "bytes": by[i], "launches": int(n[i])} for i, k in enumerate(names)}

    # ------------------------------------------------------------------ unit-test hooks
    def debug_gemm(self, A, W, *, w_type: int = 0, N: int, taps: int = 1, tap_stride: int = 0, tap_offset: int = 0,
                   l_out: Optional[int] = None, bias=None, resid=None, swiglu=False, split_k=False, simt=False):
        A = _f32(A, 3)
        B, L, C = (int(v) for v in A.shape)
        l_out = L if l_out is None else int(l_out)
        if w_type in (0, -8):
            Wb = _f32(W)
            wp = Wb.ctypes.data_as(_P)
        else:
            Wb = np.ascontiguousarray(W, dtype=np.uint8)
            wp = Wb.ctypes.data_as(_P)
        bias = _f32(bias)
        resid = _f32(resid)
        epi = (1 if bias is not None else 0) | (2 if swiglu else 0) | (4 if resid is not None else 0) | (8 if split_k else 0)
        n_out = N // 2 if swiglu else N
        out = np.empty((B, l_out, n_out), dtype=np.float32)
        st = self.lib.ace_b200_debug_gemm(self.ctx, _f(A), B, L, C, wp, int(w_type), int(N), int(taps), int(tap_stride),
                                          int(tap_offset), l_out, _f(bias), _f(resid), epi, 1 if simt else 0, _f(out),
                                          out.nbytes)
        self._ensure_ok(st, "ace_b200_debug_gemm")
        return out

    def debug_attention(self, Q, K, V, key_mask=None, n_heads: int = 1, n_kv: int = 1, window: int = -1):
        Q, K, V = _f32(Q, 3), _f32(K, 3), _f32(V, 3)
        km = _i32(key_mask)
        B, ql, kl = int(Q.shape[0]), int(Q.shape[1]), int(K.shape[1])
        out = np.empty_like(Q)
        st = self.lib.ace_b200_debug_attention(self.ctx, _f(Q), _f(K), _f(V), _i(km), B, ql, kl, int(n_heads), int(n_kv),
                                               int(window), _f(out), out.nbytes)
        self._ensure_ok(st, "ace_b200_debug_attention")
        return out

    def attention_noalloc(self, Q, K, V, key_mask=None, n_heads: int = 1, n_kv: int = 1, window: int = -1):
        """same kernel and results as debug_attention; scratch = the library's static device array (no cudaMalloc)"""
        Q, K, V = _f32(Q, 3), _f32(K, 3), _f32(V, 3)
        km = _i32(key_mask)
        B, ql, kl = int(Q.shape[0]), int(Q.shape[1]), int(K.shape[1])
        out = np.empty_like(Q)
        st = self.lib.ace_b200_attention_noalloc(self.ctx, _f(Q), _f(K), _f(V), _i(km), B, ql, kl, int(n_heads), int(n_kv),
                                                 int(window), _f(out), out.nbytes)
        self._ensure_ok(st, "ace_b200_attention_noalloc")
        return out


NOALLOC_SCRATCH_BYTES = 16 << 20  # include/acestep_b200.h: ACE_B200_NOALLOC_SCRATCH_BYTES


def pack_roundtrip(W, w_type: int, N: int, K: int, lib: Optional[ctypes.CDLL] = None):
    """CPU-only: (kernel-decoder emulation of the packed tiles, ggml dequantisation) of the same rows."""
    lib = lib or load_library()
    if w_type == 0:
        Wb = _f32(W)
    else:
        Wb = np.ascontiguousarray(W, dtype=np.uint8)
    un = np.empty((N, K), dtype=np.float32)
    ref = np.empty((N, K), dtype=np.float32)
    st = lib.ace_b200_debug_pack_roundtrip(Wb.ctypes.data_as(_P), int(w_type), int(N), int(K), _f(un), _f(ref))
    if st != 0:
        raise RuntimeError(f"ace_b200_debug_pack_roundtrip failed: {st}")
    return un, ref
