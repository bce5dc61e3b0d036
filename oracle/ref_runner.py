#!/usr/bin/env python
"""TEST INFRASTRUCTURE (oracle). Drives the UNMODIFIED reference library
(oracle/_ref/<isa>/libacestep_ggml.so) through its C ABI (acestep_ggml.h:37-108).

Must be started with LD_PRELOAD=oracle/_ref/libalignshim.so (see alignshim.c);
use oracle.refbridge.run_ref() which does that. A JSON job on argv[1] names the
model dirs, an input .npz and the calls to make; outputs are written to an .npz.

Job keys:
  lib, dit_dir?, vae_dir?, threads, compute_buffer_bytes, inputs (npz path), out (npz path)
  calls: list of
    {"op":"dit_forward","t":..,"t_r":..,"out":"name", "hidden":"hidden_states", ...}
    {"op":"dit_sample","steps":[t0,t1,...],"out":"x_final","trace":"x_steps"}   (Euler loop, acestep_ggml.cpp:2056-2086)
    {"op":"vae_decode","latents":"name","out":"audio"}
    {"op":"vae_decode_tiled","latents":"name","out":"audio","chunk":128,"overlap":32}  (acestep_ggml.cpp:2114-2215)
"""
import ctypes
import json
import sys
import time

import numpy as np


class InitParams(ctypes.Structure):
    _fields_ = [("n_threads", ctypes.c_int32), ("use_metal", ctypes.c_int32),
                ("compute_buffer_bytes", ctypes.c_size_t)]


def fptr(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float)) if a is not None else None


def iptr(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)) if a is not None else None


def bind(lib):
    F, I, P = ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_int32), ctypes.c_void_p
    lib.ace_ggml_create.argtypes = [ctypes.POINTER(InitParams), ctypes.POINTER(P)]
    lib.ace_ggml_destroy.argtypes = [P]
    lib.ace_ggml_last_error.argtypes = [P]
    lib.ace_ggml_last_error.restype = ctypes.c_char_p
    lib.ace_ggml_load_dit.argtypes = [P, ctypes.c_char_p]
    lib.ace_ggml_load_vae.argtypes = [P, ctypes.c_char_p]
    lib.ace_ggml_vae_get_info.argtypes = [P, I, I, I]
    lib.ace_ggml_vae_decode.argtypes = [P, F, ctypes.c_int32, F, ctypes.c_size_t]
    lib.ace_ggml_dit_forward.argtypes = [P, F, F, F, I, I, ctypes.c_int32, ctypes.c_int32,
                                         ctypes.c_float, ctypes.c_float, F, ctypes.c_size_t]
    i32, f32, sz = ctypes.c_int32, ctypes.c_float, ctypes.c_size_t
    lib.ace_ggml_load_text_encoder.argtypes = [P, ctypes.c_char_p]
    lib.ace_ggml_vae_encode.argtypes = [P, F, i32, F, sz]
    lib.ace_ggml_text_encoder_forward.argtypes = [P, I, i32, F, sz]
    lib.ace_ggml_text_encoder_forward_masked.argtypes = [P, I, I, i32, F, sz]
    lib.ace_ggml_text_encoder_forward_embeddings.argtypes = [P, I, i32, F, sz]
    lib.ace_ggml_text_encoder_forward_layers.argtypes = [P, I, I, i32, i32, i32, F, sz]
    lib.ace_ggml_generate_audio_simple.argtypes = [P, I, i32, i32, f32, i32, F, sz, I, I]
    lib.ace_ggml_generate_audio_style_lyric_simple.argtypes = [P, I, i32, I, i32, i32, f32, i32, F, sz, I, I]
    lib.ace_ggml_generate_audio_style_lyric_timbre_simple.argtypes = [P, I, i32, I, i32, F, I, i32, i32, i32, f32, i32, F, sz, I, I]


def main():
    job = json.load(open(sys.argv[1]))
    lib = ctypes.CDLL(job["lib"])
    bind(lib)
    ctx = ctypes.c_void_p()
    prm = InitParams(int(job.get("threads", 4)), 0, int(job.get("compute_buffer_bytes", 1 << 30)))
    assert lib.ace_ggml_create(ctypes.byref(prm), ctypes.byref(ctx)) == 0

    def check(rc, what):
        if rc != 0:
            raise RuntimeError(f"{what}: rc={rc} {lib.ace_ggml_last_error(ctx).decode()}")

    timings = {}
    t0 = time.perf_counter()
    if job.get("dit_dir"):
        check(lib.ace_ggml_load_dit(ctx, job["dit_dir"].encode()), "load_dit")
    if job.get("vae_dir"):
        check(lib.ace_ggml_load_vae(ctx, job["vae_dir"].encode()), "load_vae")
    if job.get("text_dir"):
        check(lib.ace_ggml_load_text_encoder(ctx, job["text_dir"].encode()), "load_text_encoder")
    timings["load_s"] = time.perf_counter() - t0

    inp = dict(np.load(job["inputs"])) if job.get("inputs") else {}
    outs = {}

    def get(name):
        if name is None:
            return None
        a = inp.get(name, outs.get(name))
        return None if a is None else np.ascontiguousarray(a)

    def dit_forward(c, hidden, t, t_r):
        ctxl = get(c.get("context", "context_latents"))
        enc = get(c.get("enc", "encoder_hidden_states"))
        am = get(c.get("mask")) if c.get("mask") else None
        em = get(c.get("enc_mask")) if c.get("enc_mask") else None
        T = hidden.shape[0]
        E = 0 if enc is None else enc.shape[0]
        out = np.empty((T, hidden.shape[1]), dtype=np.float32)
        check(lib.ace_ggml_dit_forward(ctx, fptr(hidden), fptr(ctxl), fptr(enc) if E else None,
                                       iptr(am), iptr(em), T, E, float(t), float(t_r),
                                       fptr(out), out.nbytes), "dit_forward")
        return out

    def vae_info():
        a, b, c = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int32()
        check(lib.ace_ggml_vae_get_info(ctx, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c)), "vae_get_info")
        return a.value, b.value, c.value

    def vae_decode(lat):
        _, ac, hop = vae_info()
        out = np.empty((lat.shape[0] * hop, ac), dtype=np.float32)
        check(lib.ace_ggml_vae_decode(ctx, fptr(lat), lat.shape[0], fptr(out), out.nbytes), "vae_decode")
        return out

    for c in job["calls"]:
        t1 = time.perf_counter()
        op = c["op"]
        if op == "dit_forward":
            outs[c["out"]] = dit_forward(c, get(c.get("hidden", "hidden_states")), c["t"], c.get("t_r", c["t"]))
        elif op == "dit_sample":
            x = get(c.get("hidden", "hidden_states")).copy()
            sched = c["steps"]
            trace = []
            step_s = []
            for i, t in enumerate(sched):
                ts = time.perf_counter()
                v = dit_forward(c, x, t, t)
                step_s.append(time.perf_counter() - ts)
                dt = t if i == len(sched) - 1 else (t - sched[i + 1])
                x = (x - v * np.float32(dt)).astype(np.float32)
                trace.append(x.copy())
            outs[c["out"]] = x
            if c.get("trace"):
                outs[c["trace"]] = np.stack(trace)
            timings[c["out"] + "_step_s"] = step_s
        elif op == "vae_decode":
            outs[c["out"]] = vae_decode(get(c["latents"]))
        elif op == "vae_decode_tiled":
            lat = get(c["latents"])
            _, ac, hop = vae_info()
            T = lat.shape[0]
            chunk, ov = int(c.get("chunk", 128)), int(c.get("overlap", 32))
            stride = max(1, chunk - 2 * ov)
            parts = []
            for core0 in range(0, T, stride):
                core1 = min(core0 + stride, T)
                w0, w1 = max(0, core0 - ov), min(T, core1 + ov)
                a = vae_decode(lat[w0:w1])
                s0 = (core0 - w0) * hop
                parts.append(a[s0: s0 + (core1 - core0) * hop])
            outs[c["out"]] = np.concatenate(parts, axis=0)
        elif op == "vae_encode":
            audio = get(c["audio"])
            lc, _, hop = vae_info()
            out = np.zeros((audio.shape[0] // hop + 2, lc), dtype=np.float32)
            check(lib.ace_ggml_vae_encode(ctx, fptr(audio), audio.shape[0], fptr(out), out.nbytes), "vae_encode")
            outs[c["out"]] = out[: audio.shape[0] // hop]
        elif op == "text_encoder":
            ids = get(c["ids"]).astype(np.int32)
            am = get(c.get("mask")) if c.get("mask") else None
            out = np.zeros((ids.shape[0], int(c["hidden"])), dtype=np.float32)
            mode = c.get("mode", "forward")
            if mode == "forward":
                rc = lib.ace_ggml_text_encoder_forward(ctx, iptr(ids), ids.shape[0], fptr(out), out.nbytes)
            elif mode == "masked":
                rc = lib.ace_ggml_text_encoder_forward_masked(ctx, iptr(ids), iptr(am), ids.shape[0], fptr(out), out.nbytes)
            elif mode == "embeddings":
                rc = lib.ace_ggml_text_encoder_forward_embeddings(ctx, iptr(ids), ids.shape[0], fptr(out), out.nbytes)
            else:
                rc = lib.ace_ggml_text_encoder_forward_layers(ctx, iptr(ids), iptr(am), ids.shape[0], int(c["n_layers"]),
                                                              int(c.get("final_norm", 1)), fptr(out), out.nbytes)
            check(rc, "text_encoder " + mode)
            outs[c["out"]] = out
        elif op == "generate":
            _, ac, hop = vae_info()
            seq_len = int(c["seq_len"])
            out = np.zeros((seq_len * hop, ac), dtype=np.float32)
            ns, nc = ctypes.c_int32(), ctypes.c_int32()
            style = get(c.get("style")) if c.get("style") else None
            lyric = get(c.get("lyric")) if c.get("lyric") else None
            refer = get(c.get("refer")) if c.get("refer") else None
            shift, seed = float(c.get("shift", 3.0)), int(c.get("seed", 0))
            if c.get("mode", "simple") == "simple":
                rc = lib.ace_ggml_generate_audio_simple(ctx, iptr(style), style.shape[0], seq_len, shift, seed, fptr(out), out.nbytes,
                                                        ctypes.byref(ns), ctypes.byref(nc))
            elif refer is None:
                rc = lib.ace_ggml_generate_audio_style_lyric_simple(ctx, iptr(style), 0 if style is None else style.shape[0], iptr(lyric),
                                                                    0 if lyric is None else lyric.shape[0], seq_len, shift, seed, fptr(out),
                                                                    out.nbytes, ctypes.byref(ns), ctypes.byref(nc))
            else:
                order = np.zeros((refer.shape[0],), np.int32)
                rc = lib.ace_ggml_generate_audio_style_lyric_timbre_simple(ctx, iptr(style), 0 if style is None else style.shape[0], iptr(lyric),
                                                                           0 if lyric is None else lyric.shape[0], fptr(refer), iptr(order),
                                                                           refer.shape[0], refer.shape[1], seq_len, shift, seed, fptr(out), out.nbytes,
                                                                           ctypes.byref(ns), ctypes.byref(nc))
            check(rc, "generate")
            outs[c["out"]] = out[: ns.value]
        else:
            raise ValueError(op)
        timings[op + "_" + c["out"] + "_s"] = time.perf_counter() - t1

    lib.ace_ggml_destroy(ctx)
    np.savez(job["out"], **outs)
    json.dump(timings, open(job["out"] + ".timings.json", "w"))


if __name__ == "__main__":
    main()
