#!/usr/bin/env python
"""Benchmark of the hot path: DiT denoise steps/s (+ seconds per 240 s song) on B200, next to the reference's
ggml CPU path on the same box.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Workload (BASELINE.json configs[1]): turbo DiT (hidden 2048, 24 layers, 16/8 heads x 128, MLP 6144), F16 GGUF,
60 s clip (T = 1500 latent frames, S = 750 tokens), E = 512 encoder tokens, batch 1 per GPU, Euler steps with the
shift-3 schedule; random-init weights (seed 0) and synthetic inputs (seed 1234) — there are no checkpoints offline.
One "step" = one Euler step = one evaluation of the velocity network for every sample of the batch.
N > 1: every rank runs its own song (independent samples, no collective in the data path) -> weak scaling.

Printed keys (one JSON line, rank 0): see the task contract; `value` is device-resident (CUDA events on the
library's stream around the K steps), `e2e` goes through the reference-facing C ABI `ace_ggml_dit_forward` with
pinned HOST buffers every step (H2D + D2H inside the timed region), `roofline` is the tap-GEMM family (tensor bound)
from a separate event-instrumented pass, `cpu_baseline` is the unmodified reference (oracle/_ref) on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ace-step-1.5-ggml_b200")]

SECONDS = 60.0
T_FRAMES = int(SECONDS * 25)
ENC_LEN = 512
QUANT = "F16"
WORKLOAD = "turbo DiT F16, 60 s clip (T=1500, S=750), E=512, batch 1 per GPU, Euler shift-3 (BASELINE configs[1])"
METRIC = "dit_denoise_steps_per_s"
UNIT = "steps/s"


def fixtures_root():
    r = os.environ.get("ACE_B200_FIXTURES", "/tmp/ace_b200_fixtures")
    os.makedirs(r, exist_ok=True)
    return r


def ensure_dit_fixture(quant=QUANT):
    from acestep_b200 import fixtures as fx
    d = os.path.join(fixtures_root(), f"turbo_{quant}")
    if not os.path.exists(os.path.join(d, "model.gguf")):
        tmp = d + f".tmp{os.getpid()}"
        lib = None
        if quant != "F16":
            from oracle import refbridge
            lib = refbridge.ref_lib()
        fx.make_dit_fixture(tmp, fx.TURBO, 0, safetensors=False, gguf_quant=quant, ggml_lib=lib, fast=True)
        try:
            os.rename(tmp, d)
        except OSError:
            pass
    return d


def ensure_vae_fixture():
    from acestep_b200 import fixtures as fx
    d = os.path.join(fixtures_root(), "vae_full")
    if not os.path.exists(os.path.join(d, "config.json")):
        fx.make_vae_fixture(d, fx.VAE_FULL, 0, fast=True)
    return d


def synth_inputs(T, E, B=1, seed=1234):
    g = np.random.default_rng(seed)
    noise = g.standard_normal((B, T, 64), dtype=np.float32)
    ctx = np.ones((B, T, 128), dtype=np.float32)
    ctx[:, :, :64] = g.standard_normal((B, T, 64), dtype=np.float32)
    enc = g.standard_normal((B, E, 2048), dtype=np.float32)
    return noise, ctx, enc


def shift_schedule(shift, n):
    u = 1.0 - np.arange(n, dtype=np.float64) / n
    return (shift * u / (1.0 + (shift - 1.0) * u)).astype(np.float32)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                pass
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx = max(mx, float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


_REAL_STDOUT = None


def guard_stdout():
    """stdout carries exactly ONE JSON line: everything libraries print (NCCL's version banner, torchrun notes) goes to stderr."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    data = (json.dumps(line) + "\n").encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def dist_setup(n_gpus):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist_mod
        torch.cuda.set_device(local)
        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = dist_mod
    return world, rank, local, dist


def run_reference_arm(a, world, rank):
    """the UNMODIFIED reference (oracle/_ref, ggml CPU backend) on this box's host cores, same workload."""
    if rank != 0:
        return
    from oracle import refbridge
    if not refbridge.ref_available():
        emit({"impl": "reference", "unavailable": "oracle/_ref not built (needs /root/reference at build time)"})
        return
    d = ensure_dit_fixture()
    noise, ctx, enc = synth_inputs(T_FRAMES, ENC_LEN)
    cores = os.cpu_count() or 1
    # bounded: one step = one full evaluation of the same network (~13 s on 16 cores), so at most 12 timed + 1 warm-up
    a.steps, a.warmup = max(1, min(a.steps, 12)), max(0, min(a.warmup, 1))
    sched = shift_schedule(3.0, max(a.steps + a.warmup, 1))
    calls = [{"op": "dit_forward", "t": float(sched[i]), "t_r": float(sched[i]), "out": f"v{i}"} for i in range(a.warmup + a.steps)]
    inp = {"hidden_states": noise[0], "context_latents": ctx[0], "encoder_hidden_states": enc[0]}
    buf = int(float(os.environ.get("ACE_REF_BUFFER_GB", "20")) * (1 << 30))
    outs, tim = refbridge.run_ref(calls, inp, dit_dir=d, threads=cores, compute_buffer_bytes=buf, env={"ACE_GGML_DIT_PROFILE": "1"}, timeout=6000)
    per = [tim[f"dit_forward_v{i}_s"] for i in range(a.warmup, a.warmup + a.steps)]
    total = float(sum(per))
    value = a.steps / total
    comp = [int(l.split("compute_ms=")[1].split()[0]) for l in tim["stderr"].splitlines() if "compute_ms=" in l]
    sample = (f"{a.steps} full DiT evaluations (T={T_FRAMES}, E={ENC_LEN}, F16) through ace_ggml_dit_forward, {cores} threads, "
              f"compute buffer {buf >> 30} GiB; wall per call {total / a.steps:.2f}s of which ggml compute_ms~{(np.mean(comp[-a.steps:]) if comp else 0):.0f}")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": 1e3 * total / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16",
            "data": "synthetic", "config": {"workload": WORKLOAD, "timing": "wall clock around each reference call"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    emit(line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-song", action="store_true", help="skip the 240 s song (DiT + VAE) measurement")
    a = ap.parse_args()
    guard_stdout()
    a.warmup = max(a.warmup, 3) if a.impl == "b200" else a.warmup
    world, rank, local, dist = dist_setup(a.gpus)

    if a.impl == "reference":
        run_reference_arm(a, world, rank)
        if dist:
            dist.destroy_process_group()
        return

    import torch
    os.environ["ACE_B200_DEVICE"] = str(local)
    torch.cuda.set_device(local)
    from acestep_b200.bridge import B200Bridge

    if rank == 0:
        ensure_dit_fixture()
    if dist:
        dist.barrier()
    d = ensure_dit_fixture()
    br = B200Bridge()  # raises if the CUDA library is missing: there is no fallback
    br.load_dit(d)
    noise, ctx, enc = synth_inputs(T_FRAMES, ENC_LEN, seed=1234 + rank)

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- device-resident: W warm-up steps, then exactly K timed steps
    sampler = ClockSampler(local)  # runs from the warm-up to the end of the e2e region (nvidia-smi needs ~100 ms per sample)
    sampler.start()
    br.generate_latents(noise, ctx, enc, schedule=shift_schedule(3.0, a.warmup))
    barrier()
    n0 = br.kernel_launches()
    t0 = time.perf_counter()
    br.generate_latents(noise, ctx, enc, schedule=shift_schedule(3.0, a.steps))
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    dev_ms = br.last_device_ms()
    launches = br.kernel_launches() - n0

    # ---- end to end through the reference-facing ABI, pinned host buffers, H2D + D2H every step
    hs = torch.from_numpy(noise[0].copy()).pin_memory().numpy()
    cx = torch.from_numpy(ctx[0]).pin_memory().numpy()
    en = torch.from_numpy(enc[0]).pin_memory().numpy()
    sched = shift_schedule(3.0, a.steps)
    for i in range(3):
        br.dit_forward_tfirst(hs, cx, en, None, None, 1.0, 1.0)
    barrier()
    t0 = time.perf_counter()
    x = hs
    for i in range(a.steps):
        v = br.dit_forward_tfirst(x, cx, en, None, None, float(sched[i]), float(sched[i]))
        dt = sched[i] if i == a.steps - 1 else sched[i] - sched[i + 1]
        x = x - v * np.float32(dt)
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop()
    h2d = int(hs.nbytes + cx.nbytes + en.nbytes)
    d2h = int(hs.nbytes)

    # ---- max over ranks
    if dist:
        tt = torch.tensor([dev_ms, e2e_ms, wall_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dev_ms, e2e_ms, wall_ms = (float(v) for v in tt.tolist())
        lt = torch.tensor([launches], device="cuda", dtype=torch.int64)
        dist.all_reduce(lt)
        launches = int(lt.item())

    if rank == 0:
        # ---- roofline of the dominant family (tap-GEMM, tensor bound), from a separate instrumented pass
        br.profile_enable(True)
        br.generate_latents(noise, ctx, enc, schedule=shift_schedule(3.0, a.steps))
        prof = br.profile_collect()
        br.profile_enable(False)
        peaks, peak_kind = measured_peaks()
        gm = prof["gemm"]
        ach = gm["flops"] / (gm["ms"] * 1e-3) / 1e12 if gm["ms"] > 0 else 0.0
        peak = float(peaks.get("bf16_tflops_sustained", 1400.0))
        total_prof = sum(v["ms"] for v in prof.values())
        traffic = None
        try:  # dram__bytes_read.sum + dram__bytes_write.sum per launch of this kernel family from the committed ncu --set full capture
            traffic = json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json")))["gemm_bytes_per_launch"]
        except Exception:
            pass
        roofline = {"bound": "tensor", "kernel": "gemm_tc_kernel (tcgen05 tap-GEMM family)", "achieved": ach, "peak": peak, "unit": "TFLOP/s",
                    "frac": ach / peak, "traffic": traffic, "peak_source": f"{peak_kind}: bf16_tflops_sustained (kernel timed inside a long step)",
                    "avg_launch_us": 1e3 * gm["ms"] / max(gm["launches"], 1), "launches": gm["launches"],
                    "flops_per_launch": gm["flops"] / max(gm["launches"], 1), "share_of_step": gm["ms"] / total_prof if total_prof else None,
                    "families_ms": {k: v["ms"] for k, v in prof.items()}}
        line = {"metric": METRIC, "value": world * a.steps / (dev_ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": dev_ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16",
                "data": "synthetic", "config": {"workload": WORKLOAD, "global_batch": world, "parallelism": f"independent samples x{world}",
                                                 "l2": "weights (3.0 GB F16) exceed L2 and are re-streamed every step; no explicit flush",
                                                 "timing": "CUDA events on the library stream around the K steps, max over ranks"},
                "wall_ms_per_step": wall_ms / a.steps, "clocks": clocks, "gpu_launches": launches,
                "e2e": {"value": world * a.steps / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "api": "ace_ggml_dit_forward per step, pinned host buffers, host Euler update"},
                "roofline": roofline}
        if world == 1 and not a.no_song:
            try:
                line["song_240s"] = measure_song(br)
            except Exception as ex:  # a measurement extra must not lose the headline line
                line["song_240s"] = {"error": str(ex)[:200]}
        if world == 1 and not a.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(d, noise, ctx, enc)
        emit(line)
    br.close()
    if dist:
        dist.barrier()
        dist.destroy_process_group()


def measure_song(br):
    """seconds per 240 s song: 8-step turbo DiT + VAE decode, batch 1, device-resident (one H2D, one D2H)."""
    vd = ensure_vae_fixture()
    br.load_vae(vd)
    T = 240 * 25
    noise, ctx, enc = synth_inputs(T, ENC_LEN, seed=99)
    sched = shift_schedule(3.0, 8)
    audio = np.empty((1, T * br.hop_length, br.audio_channels), np.float32)
    br.generate_audio(noise, ctx, enc, schedule=sched, out_audio=audio)
    t0 = time.perf_counter()
    br.generate_audio(noise, ctx, enc, schedule=sched, out_audio=audio)
    wall = time.perf_counter() - t0
    return {"value": br.last_device_ms() * 1e-3, "unit": "s (device, 8 DiT steps + VAE decode)", "wall_s_incl_copies": wall,
            "audio_seconds": 240, "finite": bool(np.isfinite(audio).all())}


def cpu_baseline(d, noise, ctx, enc):
    from oracle import refbridge
    if not refbridge.ref_available():
        return {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}
    cores = os.cpu_count() or 1
    buf = int(float(os.environ.get("ACE_REF_BUFFER_GB", "20")) * (1 << 30))
    inp = {"hidden_states": noise[0], "context_latents": ctx[0], "encoder_hidden_states": enc[0]}
    calls = [{"op": "dit_forward", "t": 1.0, "t_r": 1.0, "out": "v0"}, {"op": "dit_forward", "t": 0.9, "t_r": 0.9, "out": "v1"}]
    outs, tim = refbridge.run_ref(calls, inp, dit_dir=d, threads=cores, compute_buffer_bytes=buf, env={"ACE_GGML_DIT_PROFILE": "1"}, timeout=3000)
    s = tim["dit_forward_v1_s"]
    comp = [l.split("compute_ms=")[1].split()[0] for l in tim["stderr"].splitlines() if "compute_ms=" in l]
    return {"value": 1.0 / s, "unit": UNIT, "cores": cores, "kind": "reference",
            "sample": f"1 of the 8 evaluations of the same workload (2nd call; wall {s:.2f}s, ggml compute_ms={comp[-1] if comp else '?'}, buffer {buf >> 30} GiB)"}


if __name__ == "__main__":
    main()
