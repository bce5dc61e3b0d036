#!/usr/bin/env python
"""Benchmark of the hot path — DiT denoise steps/s and seconds per 240 s song on B200 — next to the reference's ggml CPU
path on the same box.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload config2|config1|base50|vae600]

Default workload = BASELINE.json configs[2], the configuration the metric is quoted on:
  turbo DiT (hidden 2048, 24 layers, 16/8 heads x 128, MLP 6144), Q4_K GGUF, 240 s clip (T = 6000 latent frames, S = 3000 tokens),
  E = 512 encoder tokens, a FIXED batch of 8 songs x CFG (guidance 7, unconditional branch = all-zero encoder states) = 16 network
  evaluations per step, dealt over the N ranks by song (both CFG branches of a song stay on one rank; no collective in the data
  path) -> STRONG scaling; random-init weights (seed 0) and synthetic inputs (seeds 1234..1241) — there are no checkpoints offline.
  One "step" = one Euler step of the WHOLE batch.  `song_240s` = 8 such steps + the VAE decode of the 8 songs.
Other workloads: config1 (F16, 60 s, one song per GPU, weak scaling: round 1's line), base50 (configs[4]: Q8_0, 120 s, 50-step closed-form
schedule, 8 songs x CFG + VAE), vae600 (configs[3]: VAE decode of one 600 s clip; N > 1 = time-split with a 16-frame halo, core slices
written straight into rank 0's output buffer over NVLink).

Keys of the JSON line (rank 0): the task contract's.  `value` = device-resident (CUDA events on the library's stream around the K
steps, inputs already in HBM), max over ranks; `e2e` = the same steps through the C ABI with pinned HOST buffers every step (H2D of the
2B-evaluation batch + D2H of its velocities inside the timed region, CFG combine and Euler update on the host); `roofline` = tap-GEMM
family (tensor bound) from a separate event-instrumented pass, plus the element-wise family's HBM fraction; `parity` = this run's GPU
outputs against the reference's on the same inputs (bench fails when it is out of tolerance); `cpu_baseline` / `--impl reference` = the
unmodified reference (oracle/_ref) on the host cores: a BOUNDED sample of the same workload — one evaluation at S = 3000 truncated to 1
and 2 layers (ACE_GGML_DIT_MAX_LAYERS; the full depth needs ~90 GB of score tensors on the CPU), extrapolated linearly in depth to the
24-layer, 16-evaluation step and stated as such.
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
sys.path[:0] = [ROOT, os.path.join(ROOT, "ace-step-1.5-ggml_b200"), os.path.join(ROOT, "tests")]

ENC_LEN = 512
GUIDANCE = 7.0
METRIC = "dit_denoise_steps_per_s"
UNIT = "steps/s"
WORKLOADS = {
    "config2": dict(quant="Q4_K", seconds=240, songs=8, cfg=True, song_steps=8, scaling="strong", dtype="f16",
                    desc="turbo DiT Q4_K, 240 s clip (T=6000, S=3000), E=512, 8 songs x CFG(g=7) = 16 evaluations/step dealt over the GPUs by song, "
                         "Euler shift-3 (BASELINE configs[2])"),
    "base50": dict(quant="Q8_0", seconds=120, songs=8, cfg=True, song_steps=50, scaling="strong", dtype="f16",
                   desc="turbo-dim DiT Q8_0, 120 s clip (T=3000, S=1500), E=512, 8 songs x CFG(g=7), 50-step closed-form shift-3 schedule "
                        "(BASELINE configs[4])"),
    "config1": dict(quant="F16", seconds=60, songs=0, cfg=False, song_steps=8, scaling="weak", dtype="f16",
                    desc="turbo DiT F16, 60 s clip (T=1500, S=750), E=512, batch 1 per GPU, Euler shift-3 (BASELINE configs[1])"),
}
TOL = {"F16": (5e-3, 0.99999), "Q8_0": (2e-2, 0.9998), "Q6_K": (2e-2, 0.9998), "Q4_K": (2e-2, 0.9998)}


def synth_inputs(T, E, B=1, seed=1234):
    """per song s: the inputs tools.fixtures.dit_inputs(TURBO, T, E, seed + s) (what the parity goldens use)"""
    from tools import fixtures as fx
    ins = [fx.dit_inputs(fx.TURBO, T, E, seed=seed + b) for b in range(B)]
    return (np.stack([i["hidden_states"] for i in ins]), np.stack([i["context_latents"] for i in ins]),
            np.stack([i["encoder_hidden_states"] for i in ins]))


def shift_schedule(shift, n):
    u = 1.0 - np.arange(n, dtype=np.float64) / n
    return (shift * u / (1.0 + (shift - 1.0) * u)).astype(np.float32)


def metrics(a, ref):
    a = np.asarray(a, np.float64).ravel()
    r = np.asarray(ref, np.float64).ravel()
    d = a - r
    return {"rel_l2": float(np.linalg.norm(d) / (np.linalg.norm(r) + 1e-30)),
            "cos": float(a @ r / (np.linalg.norm(a) * np.linalg.norm(r) + 1e-30))}


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
        sm, pw, mx, reasons = [], [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx = max(mx, float(f[1]))
                pw.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm),
                "power_w_max": max(pw) if pw else None}


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback (B200_PROFILING.md)"


_REAL_STDOUT = None


def guard_stdout():
    """stdout carries exactly ONE JSON line: everything libraries print (NCCL's banner, torchrun notes) goes to stderr."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, (json.dumps(line) + "\n").encode())


def log(*a):
    print("[bench]", *a, file=sys.stderr, flush=True)


def dist_setup():
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


# ------------------------------------------------------------------------------------------------ reference (CPU) legs

def ref_threads():
    return os.cpu_count() or 1


def reference_sample(wl, layers=(1, 2), want_outputs=False):
    """Bounded sample of the workload on the unmodified reference: ONE network evaluation of song 0's conditional branch at the
    workload's sequence length, truncated to 1 and 2 layers.  Returns per-evaluation seconds extrapolated to the full depth
    (fixed part + 24 x per-layer part), from wall time and from ggml's own compute_ms, plus the raw outputs."""
    from oracle import refbridge
    from parity_common import turbo_dit_dir
    d = turbo_dit_dir(wl["quant"])
    T = int(wl["seconds"] * 25)
    hs, cx, en = synth_inputs(T, ENC_LEN, 1)
    inp = {"hidden_states": hs[0], "context_latents": cx[0], "encoder_hidden_states": en[0]}
    S = (T + 1) // 2
    wall, comp, outs = {}, {}, {}
    bufs = {}
    for k in layers:
        # the reference allocates every intermediate from the arena (no reuse) and zero-fills it on every call
        # (acestep_ggml.cpp:1351): [S,S,16] F32 scores x ~5 copies per self-attention dominate
        buf = int((3.0 + k * (S * S * 16 * 4 * 5.5 + S * 6144 * 4 * 12) / 2 ** 30) * 2 ** 30)
        bufs[k] = buf
        o, tim = refbridge.run_ref([{"op": "dit_forward", "t": 0.7, "t_r": 0.7, "out": "v"}], inp, dit_dir=d, threads=ref_threads(), compute_buffer_bytes=buf,
                                   env={"ACE_GGML_DIT_PROFILE": "1", "ACE_GGML_DIT_MAX_LAYERS": str(k)}, timeout=3000)
        wall[k] = float(tim["dit_forward_v_s"])
        c = [float(l.split("compute_ms=")[1].split()[0]) for l in tim["stderr"].splitlines() if "compute_ms=" in l]
        comp[k] = (c[-1] if c else float("nan")) * 1e-3
        outs[k] = o["v"]
        log(f"reference sample: {wl['quant']} T={T} k={k}: wall {wall[k]:.2f}s compute {comp[k]:.2f}s arena {buf / 2 ** 30:.1f} GiB")
    k0, k1 = layers[0], layers[-1]
    n_layers = 24

    def extrap(t):
        per = (t[k1] - t[k0]) / (k1 - k0)
        return max(t[k0] - k0 * per, 0.0) + n_layers * per

    return {"eval_wall_s": extrap(wall), "eval_compute_s": extrap(comp), "wall": wall, "compute": comp, "arena_gib": {k: b / 2 ** 30 for k, b in bufs.items()},
            "outputs": outs if want_outputs else None, "inputs": inp, "dir": d, "layers": list(layers)}


def evals_per_step(wl, world=1):
    songs = wl["songs"] if wl["songs"] else world
    return songs * (2 if wl["cfg"] else 1)


def reference_record(wl, rs):
    n_ev = evals_per_step(wl)
    step_wall, step_comp = n_ev * rs["eval_wall_s"], n_ev * rs["eval_compute_s"]
    sample = (f"1 evaluation of song 0 (cond branch) at T={int(wl['seconds'] * 25)}, E={ENC_LEN}, {wl['quant']}, truncated to {rs['layers']} layers "
              f"(wall {', '.join(f'{rs['wall'][k]:.2f}' for k in rs['layers'])} s; ggml compute_ms {', '.join(f'{rs['compute'][k]:.2f}' for k in rs['layers'])} s; "
              f"arena {', '.join(f'{rs['arena_gib'][k]:.0f}' for k in rs['layers'])} GiB, zero-filled by the reference every call), extrapolated linearly in depth to 24 layers "
              f"= {rs['eval_wall_s']:.1f} s per evaluation x {n_ev} evaluations per step; full depth does not fit host RAM at S=3000")
    return {"value": 1.0 / step_wall, "unit": UNIT, "cores": ref_threads(), "kind": "reference", "sample": sample,
            "value_compute_ms_only": 1.0 / step_comp if step_comp > 0 else None, "seconds_per_step": step_wall}


def run_reference_arm(a, rank):
    """the UNMODIFIED reference (oracle/_ref, ggml CPU backend) on this box's host cores: bounded sample of the same workload."""
    if rank != 0:
        return
    from oracle import refbridge
    wl = WORKLOADS[a.workload if a.workload in WORKLOADS else "config2"]
    if not refbridge.ref_available():
        emit({"impl": "reference", "unavailable": "oracle/_ref not built (needs /root/reference at build time)"})
        return
    # every "step" of this arm is one bounded sample (two truncated evaluations); keep the run within a few minutes
    steps, warm = max(1, min(a.steps, 2)), 0
    recs = [reference_record(wl, reference_sample(wl)) for _ in range(steps)]
    best = max(recs, key=lambda r: r["value"])
    line = {"impl": "reference", "metric": METRIC, "value": best["value"], "unit": UNIT, "n_gpus": a.gpus, "steps": steps, "warmup": warm,
            "ms_per_step": 1e3 * best["seconds_per_step"], "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None, "dtype": "int8 x " + wl["quant"],
            "data": "synthetic", "config": {"workload": wl["desc"], "timing": "wall clock around each reference call, best of the samples"},
            "cpu_baseline": best, "e2e": {"value": best["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    emit(line)


# ------------------------------------------------------------------------------------------------ B200 arm

def pinned(a):
    import torch
    t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    return t.numpy(), t  # keep the tensor alive


def run_dit_workload(a, wl, world, rank, local, dist):
    import torch
    from acestep_b200 import shard
    from acestep_b200.bridge import B200Bridge
    from parity_common import full_vae_dir, turbo_dit_dir

    T = int(wl["seconds"] * 25)
    n_songs = wl["songs"] if wl["songs"] else world
    mine = shard.song_indices(n_songs, rank, world)
    if rank == 0:
        turbo_dit_dir(wl["quant"])
        full_vae_dir()
    if dist:
        dist.barrier()
    d = turbo_dit_dir(wl["quant"])
    br = B200Bridge()  # raises if the CUDA library is missing: there is no fallback
    br.load_dit(d)
    B = len(mine)
    hs_all, cx_all, en_all = [], [], []
    for s in mine:
        h, c, e = synth_inputs(T, ENC_LEN, 1, seed=1234 + s)
        hs_all.append(h[0]); cx_all.append(c[0]); en_all.append(e[0])
    noise, ctx, enc = np.stack(hs_all), np.stack(cx_all), np.stack(en_all)
    unc = np.zeros_like(enc) if wl["cfg"] else None
    g = GUIDANCE if wl["cfg"] else 1.0

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
            torch.cuda.synchronize()

    def sample(k):
        return br.generate_latents(noise, ctx, enc, uncond_enc=unc, schedule=shift_schedule(3.0, k), guidance=g)

    # ---- device-resident: W warm-up steps, then exactly K timed steps
    sampler = ClockSampler(local)
    sampler.start()
    sample(a.warmup)
    barrier()
    n0 = br.kernel_launches()
    t0 = time.perf_counter()
    sample(a.steps)
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    dev_ms = br.last_device_ms()
    launches = br.kernel_launches() - n0

    # ---- end to end through the C ABI: pinned host buffers, H2D of the whole evaluation batch + D2H of its velocities every step
    EB = B * (2 if wl["cfg"] else 1)
    hid2, _k1 = pinned(np.concatenate([noise, noise]) if wl["cfg"] else noise)
    ctx2, _k2 = pinned(np.concatenate([ctx, ctx]) if wl["cfg"] else ctx)
    enc2, _k3 = pinned(np.concatenate([enc, unc]) if wl["cfg"] else enc)
    vout, _k4 = pinned(np.empty_like(hid2))
    sched = shift_schedule(3.0, a.steps)
    one = np.ones((EB,), np.float32)
    for i in range(2):
        br.dit_forward_batched(hid2, ctx2, enc2, None, None, one, one, out=vout)
    barrier()
    # host side of a step (what the caller's pipeline does between two library calls): CFG combine + Euler update, on torch CPU views
    # of the pinned buffers (multi-threaded, in place)
    t_hid, t_v = torch.from_numpy(hid2), torch.from_numpy(vout)
    t_x = torch.from_numpy(noise.copy())
    t_tmp = torch.empty_like(t_x)
    t0 = time.perf_counter()
    for i in range(a.steps):
        t_hid[:B].copy_(t_x)
        if wl["cfg"]:
            t_hid[B:].copy_(t_x)
        t = float(sched[i])
        br.dit_forward_batched(hid2, ctx2, enc2, None, None, one * t, one * t, out=vout)
        dt = float(sched[i] if i == a.steps - 1 else sched[i] - sched[i + 1])
        if wl["cfg"]:  # v = v_u + g (v_c - v_u)
            torch.sub(t_v[:B], t_v[B:], out=t_tmp)
            t_tmp.mul_(g).add_(t_v[B:])
            t_x.sub_(t_tmp, alpha=dt)
        else:
            t_x.sub_(t_v, alpha=dt)
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop()
    h2d = int(hid2.nbytes + ctx2.nbytes + enc2.nbytes)
    d2h = int(vout.nbytes)

    # ---- the song: song_steps Euler steps + VAE decode of this rank's songs, latents never leave HBM
    song = None
    if not a.no_song:
        br.load_vae(full_vae_dir())
        audio = np.empty((B, T * br.hop_length, br.audio_channels), np.float32)
        ssched = shift_schedule(3.0, wl["song_steps"])
        br.generate_audio(noise, ctx, enc, uncond_enc=unc, schedule=ssched, guidance=g, out_audio=audio)
        barrier()
        t0 = time.perf_counter()
        br.generate_audio(noise, ctx, enc, uncond_enc=unc, schedule=ssched, guidance=g, out_audio=audio)
        barrier()
        song = [br.last_device_ms() * 1e-3, time.perf_counter() - t0, float(np.isfinite(audio).all())]
        del audio

    # ---- max over ranks
    if dist:
        vals = [dev_ms, e2e_ms, wall_ms] + (song[:2] if song else [0.0, 0.0])
        tt = torch.tensor(vals, device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dev_ms, e2e_ms, wall_ms, s0, s1 = (float(v) for v in tt.tolist())
        lt = torch.tensor([launches, h2d, d2h, int(song[2]) if song else 1], device="cuda", dtype=torch.int64)
        dist.all_reduce(lt)
        launches, h2d, d2h = int(lt[0]), int(lt[1]), int(lt[2])
        if song:
            song = [s0, s1, float(int(lt[3]) == world)]

    ok = True
    if rank == 0:
        # ---- roofline of the dominant family (tap-GEMM, tensor bound) + the element-wise family's HBM fraction: separate instrumented pass
        br.profile_enable(True)
        sample(max(2, min(a.steps, 4)))
        prof = br.profile_collect()
        br.profile_enable(False)
        peaks, peak_kind = measured_peaks()
        gm, el, at = prof["gemm"], prof["elementwise"], prof["attention"]
        ach = gm["flops"] / (gm["ms"] * 1e-3) / 1e12 if gm["ms"] > 0 else 0.0
        peak = float(peaks.get("bf16_tflops_sustained", 1400.0))
        hbm = float(peaks.get("hbm_gbs", 6650.0))
        total_prof = sum(v["ms"] for v in prof.values())
        traffic = None
        try:  # dram bytes per launch of this kernel family from the committed ncu --set full capture of this workload
            traffic = json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json"))).get(f"gemm_bytes_per_launch_{a.workload}")
        except Exception:
            pass
        roofline = {"bound": "tensor", "kernel": "gemm_tc2_kernel / gemm_tc_kernel (tcgen05 tap-GEMM family; packed weights of launches >= 2048 rows are decoded per launch by predecode_kernel, whose time is inside the family)", "achieved": ach, "peak": peak, "unit": "TFLOP/s",
                    "frac": ach / peak, "traffic": traffic, "peak_source": f"{peak_kind}: bf16_tflops_sustained (kernel timed inside a long step)",
                    "avg_launch_us": 1e3 * gm["ms"] / max(gm["launches"], 1), "launches": gm["launches"],
                    "flops_per_launch": gm["flops"] / max(gm["launches"], 1),
                    "algorithmic_bytes_per_launch": gm["bytes"] / max(gm["launches"], 1), "share_of_step": gm["ms"] / total_prof if total_prof else None,
                    "families_ms": {k: v["ms"] for k, v in prof.items()},
                    "attention": {"bound": "tensor", "achieved": at["flops"] / (at["ms"] * 1e-3) / 1e12 if at["ms"] > 0 else 0.0, "peak": peak, "unit": "TFLOP/s",
                                  "frac": (at["flops"] / (at["ms"] * 1e-3) / 1e12 / peak) if at["ms"] > 0 else 0.0, "share_of_step": at["ms"] / total_prof if total_prof else None},
                    "elementwise": {"bound": "hbm", "achieved": el["bytes"] / (el["ms"] * 1e-3) / 1e9 if el["ms"] > 0 else 0.0, "peak": hbm, "unit": "GB/s",
                                    "frac": (el["bytes"] / (el["ms"] * 1e-3) / 1e9 / hbm) if el["ms"] > 0 else 0.0, "share_of_step": el["ms"] / total_prof if total_prof else None}}
        n_steps_total = a.steps  # one step advances the whole batch (strong) / every rank's song (weak: world songs)
        value = (world if wl["scaling"] == "weak" else 1) * n_steps_total / (dev_ms * 1e-3)
        e2e_value = (world if wl["scaling"] == "weak" else 1) * n_steps_total / (e2e_ms * 1e-3)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": dev_ms / a.steps, "higher_is_better": True, "scaling": wl["scaling"], "vs_baseline": None, "dtype": wl["dtype"],
                "data": "synthetic",
                "config": {"workload": wl["desc"], "global_batch": n_songs, "evaluations_per_step": evals_per_step(wl, world),
                           "parallelism": f"songs dealt round-robin over {world} GPU(s), both CFG branches of a song on one GPU, no collective",
                           "evaluation_batching": f"all {EB} evaluations of a rank in ONE batched forward (M = {EB * ((T + 1) // 2)} rows per GEMM)",
                           "l2": "weights (0.9-3.2 GB) and per-layer activations exceed the 126 MB L2 and are re-streamed every step; no explicit flush",
                           "timing": "CUDA events on the library stream around the K steps, max over ranks"},
                "wall_ms_per_step": wall_ms / a.steps, "clocks": clocks, "gpu_launches": launches,
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "api": "ace_b200_dit_forward_batched per step (2B-evaluation batch), pinned host buffers, CFG combine + Euler update on the host"},
                "evaluations_per_s": value * evals_per_step(wl, world) / (world if wl["scaling"] == "weak" else 1),
                "roofline": roofline}
        if song:
            line["song_240s" if wl["seconds"] == 240 else "song"] = {
                "seconds_per_batch": song[0], "seconds_per_song": song[0] / n_songs, "songs": n_songs, "audio_seconds_per_song": wl["seconds"],
                "unit": f"s (device, {wl['song_steps']} DiT steps x {evals_per_step(wl, world)} evaluations + VAE decode, max over ranks)",
                "wall_s_incl_copies": song[1], "finite": bool(song[2])}
        if world == 1 and not a.no_cpu_baseline:
            try:
                cb, par = cpu_baseline_and_parity(br, wl)
                line["cpu_baseline"], line["parity"] = cb, par
                ok = ok and par["ok"]
            except Exception as ex:  # a broken CPU leg must not lose the GPU measurement — but the run still fails
                line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": f"failed: {str(ex)[:300]}"}
                ok = False
            if a.workload == "config2" and not a.no_config1:
                try:
                    line["config1"] = config1_record(br)
                    ok = ok and line["config1"]["parity"]["ok"]
                except Exception as ex:
                    line["config1"] = {"error": str(ex)[:300]}
                    ok = False
        emit(line)
    br.close()
    if dist:
        dist.barrier()
    return ok


def cpu_baseline_and_parity(br, wl):
    """rank 0, N = 1: the reference's bounded sample of this workload on the host cores, and the GPU's outputs for the same truncated
    evaluations compared with it (parity on the bench's own configuration)."""
    from oracle import refbridge
    if not refbridge.ref_available():
        return ({"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"},
                {"ok": False, "config": "oracle/_ref not built"})
    rs = reference_sample(wl, want_outputs=True)
    inp = rs["inputs"]
    rl, cs = TOL[wl["quant"]]
    worst = {"rel_l2": 0.0, "cos": 1.0}
    prev = os.environ.get("ACE_GGML_DIT_MAX_LAYERS")
    try:
        for k in rs["layers"]:
            os.environ["ACE_GGML_DIT_MAX_LAYERS"] = str(k)
            out = br.dit_forward_tfirst(inp["hidden_states"], inp["context_latents"], inp["encoder_hidden_states"], None, None, 0.7, 0.7)
            m = metrics(out, rs["outputs"][k])
            worst = {"rel_l2": max(worst["rel_l2"], m["rel_l2"]), "cos": min(worst["cos"], m["cos"])}
    finally:
        if prev is None:
            os.environ.pop("ACE_GGML_DIT_MAX_LAYERS", None)
        else:
            os.environ["ACE_GGML_DIT_MAX_LAYERS"] = prev
    par = {"config": f"{wl['quant']} T={int(wl['seconds'] * 25)} E={ENC_LEN}, song 0 cond branch, network truncated to {rs['layers']} layers (what the CPU reference can hold); "
                     "full depth at turbo dims is pinned by tests/test_turbo_parity_gpu.py and the config1 sub-record",
           "rel_l2": worst["rel_l2"], "cos": worst["cos"], "tol": {"rel_l2": rl, "cos": cs}, "ok": bool(worst["rel_l2"] <= rl and worst["cos"] >= cs)}
    return reference_record(wl, rs), par


def config1_record(br):
    """BASELINE configs[1] (F16, 60 s, batch 1) at FULL depth on both sides: the GPU's steps/s (device-resident and through
    ace_ggml_dit_forward with host buffers) next to one full reference evaluation of the same inputs, and their parity."""
    from oracle import refbridge
    from parity_common import turbo_dit_dir
    wl = WORKLOADS["config1"]
    T = 1500
    d = turbo_dit_dir("F16")
    br.load_dit(d)
    hs, cx, en = synth_inputs(T, ENC_LEN, 1)
    K = 16
    br.generate_latents(hs, cx, en, schedule=shift_schedule(3.0, 4))
    br.generate_latents(hs, cx, en, schedule=shift_schedule(3.0, K))
    dev = K / (br.last_device_ms() * 1e-3)
    h0, _a = pinned(hs[0]); c0, _b = pinned(cx[0]); e0, _c = pinned(en[0])
    for _ in range(2):
        v_gpu = br.dit_forward_tfirst(h0, c0, e0, None, None, 1.0, 1.0)
    t0 = time.perf_counter()
    for i in range(K):
        br.dit_forward_tfirst(h0, c0, e0, None, None, 1.0, 1.0)
    e2e = K / (time.perf_counter() - t0)
    buf = int(float(os.environ.get("ACE_REF_BUFFER_GB", "16")) * (1 << 30))  # the reference needs 14.01 GiB of arena at S=750, E=512
    inp = {"hidden_states": hs[0], "context_latents": cx[0], "encoder_hidden_states": en[0]}
    calls = [{"op": "dit_forward", "t": 1.0, "t_r": 1.0, "out": "w"}, {"op": "dit_forward", "t": 1.0, "t_r": 1.0, "out": "v"}]
    outs, tim = refbridge.run_ref(calls, inp, dit_dir=d, threads=ref_threads(), compute_buffer_bytes=buf, env={"ACE_GGML_DIT_PROFILE": "1"}, timeout=3000)
    s = float(tim["dit_forward_v_s"])
    comp = [float(l.split("compute_ms=")[1].split()[0]) for l in tim["stderr"].splitlines() if "compute_ms=" in l]
    m = metrics(v_gpu, outs["v"])
    rl, cs = TOL["F16"]
    return {"workload": wl["desc"], "value": dev, "e2e": e2e, "unit": UNIT,
            "cpu_reference": {"value": 1.0 / s, "value_compute_ms_only": (1e3 / comp[-1]) if comp else None, "cores": ref_threads(), "arena_gib": buf / 2 ** 30,
                              "sample": f"one FULL evaluation (24 layers) of the same inputs, 2nd call of the process: wall {s:.2f}s, ggml compute_ms {comp[-1] if comp else '?'}"},
            "parity": {"config": "F16 T=1500 E=512 full depth, t=1.0", "rel_l2": m["rel_l2"], "cos": m["cos"], "tol": {"rel_l2": rl, "cos": cs},
                       "ok": bool(m["rel_l2"] <= rl and m["cos"] >= cs)}}


def run_vae600(a, world, rank, local, dist):
    """BASELINE configs[3]: VAE decode of ONE 600 s stereo clip (T = 15000 latent frames -> 28.8 M samples x 2).  N = 1: un-tiled decode.
    N > 1: time-split, rank r decodes its core +- 16 halo frames and its last kernel writes the core slice into rank 0's output
    buffer over NVLink (CUDA IPC peer mapping).  step = one decode of the clip; value = clips/s; also x real time."""
    import torch
    from acestep_b200 import shard
    from acestep_b200.bridge import B200Bridge
    from parity_common import full_vae_dir
    from tools import fixtures as fx
    T = 15000
    if rank == 0:
        full_vae_dir()
    if dist:
        dist.barrier()
    br = B200Bridge()
    br.load_vae(full_vae_dir())
    lat = fx.vae_latents(fx.VAE_FULL, T, seed=7)

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
            torch.cuda.synchronize()

    split = None
    if world > 1:
        split = shard.PeerSplitDecode(br, T, rank, world, dist, halo=16)
    sampler = ClockSampler(local)
    sampler.start()
    times = []
    n0 = 0
    for i in range(a.warmup + a.steps):
        if i == a.warmup:
            n0 = br.kernel_launches()
        barrier()
        if split:
            split.decode(lat)
        else:
            br.vae_decode_tfirst(lat)
        barrier()
        times.append(br.last_device_ms())
    launches = br.kernel_launches() - n0
    dev_ms = float(np.mean(times[a.warmup:]))
    # e2e: host latents in, host audio out
    audio = None
    t0 = time.perf_counter()
    for i in range(max(1, a.steps // 2)):
        if split:
            audio = split.decode(lat, fetch=True)
        else:
            audio = br.vae_decode_tfirst(lat)
        barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / max(1, a.steps // 2)
    clocks = sampler.stop()
    if dist:
        tt = torch.tensor([dev_ms, e2e_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dev_ms, e2e_ms = (float(v) for v in tt.tolist())
        lt = torch.tensor([launches], device="cuda", dtype=torch.int64)
        dist.all_reduce(lt)
        launches = int(lt[0])
    if rank == 0:
        peaks, peak_kind = measured_peaks()
        flops = 4.874e9 * T
        peak = float(peaks.get("bf16_tflops_sustained", 1400.0))
        ach = flops / (dev_ms * 1e-3) / 1e12
        line = {"metric": "vae_decode_clips_per_s", "value": 1e3 / dev_ms, "unit": "600 s clips/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": dev_ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
                "config": {"workload": "VAE decode only, 600 s clip stereo (T=15000 frames -> 28.8 M samples x 2), F16 (BASELINE configs[3])",
                           "parallelism": "un-tiled decode on one GPU" if world == 1 else f"time-split over {world} GPUs, halo 16 frames, core slices peer-written into rank 0's buffer over NVLink",
                           "l2": "activations (GBs per level) exceed L2", "timing": "CUDA events on the library stream, mean of K decodes, max over ranks"},
                "x_realtime": 600.0 / (dev_ms * 1e-3), "clocks": clocks, "gpu_launches": launches,
                "e2e": {"value": 1e3 / e2e_ms, "unit": "600 s clips/s", "h2d_bytes_per_step": int(lat.nbytes), "d2h_bytes_per_step": int(T * 1920 * 2 * 4),
                        "api": "ace_ggml_vae_decode (N=1) / PeerSplitDecode (N>1) with host latents in, host audio out"},
                "roofline": {"bound": "tensor", "kernel": "gemm_tc_kernel (tap-GEMM convolutions)", "achieved": ach * world ** 0, "peak": peak * world, "unit": "TFLOP/s",
                             "frac": ach / (peak * world), "traffic": None, "peak_source": f"{peak_kind}: bf16_tflops_sustained x {world} GPU(s)",
                             "flops_per_step": flops},
                "finite": bool(np.isfinite(audio).all()) if audio is not None else None}
        emit(line)
    br.close()
    if dist:
        dist.barrier()
    return True


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="config2", choices=["config2", "config1", "base50", "vae600"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-config1", action="store_true", help="skip the configs[1] full-depth sub-record (config2 workload, N = 1)")
    ap.add_argument("--no-song", action="store_true", help="skip the song (DiT steps + VAE decode) measurement")
    a = ap.parse_args()
    guard_stdout()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    if a.impl == "reference":
        run_reference_arm(a, rank)  # rank 0 alone runs it; no process group is needed
        return 0
    a.warmup = max(a.warmup, 3)
    a.steps = max(a.steps, 1)
    world, rank, local, dist = dist_setup()
    import torch
    os.environ["ACE_B200_DEVICE"] = str(local)
    torch.cuda.set_device(local)
    if a.workload == "vae600":
        ok = run_vae600(a, world, rank, local, dist)
    else:
        ok = run_dit_workload(a, WORKLOADS[a.workload], world, rank, local, dist)
    if dist:
        dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
