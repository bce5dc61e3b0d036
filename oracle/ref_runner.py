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
        else:
            raise ValueError(op)
        timings[op + "_" + c["out"] + "_s"] = time.perf_counter() - t1

    lib.ace_ggml_destroy(ctx)
    np.savez(job["out"], **outs)
    json.dump(timings, open(job["out"] + ".timings.json", "w"))


if __name__ == "__main__":
    main()
