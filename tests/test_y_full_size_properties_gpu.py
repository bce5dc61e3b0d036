"""Size-independent properties at BASELINE.json's FULL sizes — the shapes the CPU reference cannot run to completion (configs[2]: Q4_K,
240 s, 24 layers at S = 3000 needs ~90 GB of score tensors on the host; configs[3]: one 600 s clip; configs[4]: 50 steps x CFG), so the
GPU path is held to identities that must be true whatever the numbers are.  The reference semantics each identity restates:

  * samples of a batch are independent (acestep_ggml.cpp:1304-1482 evaluates one sample per call; ace_b200_dit_forward_batched must
    therefore equal B single calls, and a sample's result must not depend on who shares its batch);
  * CFG `v = v_u + g (v_c - v_u)` at g = 1 is the conditional velocity, and with identical branches it is that velocity
    for EVERY g — so the 2B-sample CFG sampler must reproduce the plain sampler;
  * the device-resident sampler is the host loop `x <- x - v dt` of acestep_ggml.cpp:2056-2086 around ace_ggml_dit_forward;
  * masked (padded) frames do not influence unmasked ones (build_attention_mask, acestep_dit_model.cpp:1132-1173);
  * the VAE decoder is a convolution stack with a 9.5-frame receptive field (acestep_vae_model.cpp:957-1002): a window decoded with a
    16-frame halo equals the same frames of the whole clip — what the chunk/overlap policy of acestep_ggml.cpp:2114-2229 and the
    8-GPU time split rely on — and clips of a batch are independent.

Bounds.  Where both sides run the same kernels with the same launch geometry on the same rows the comparison is BIT-EXACT (the K split
of the accumulating GEMMs is the one geometry choice that depends on the batch: ACE_B200_GEMM_SPLITK=1 pins it for those checks).
Where the geometry or the host arithmetic differs (fp32 re-association of a K split, `x - v dt` as one FMA on the device against
NumPy's multiply-then-subtract) the two runs start 1e-7 apart and every fp16 operand rounding of the 24 layers (xn, q/k/v, attention
output, SwiGLU output: ulp 4.9e-4) turns that into decorrelated rounding noise: the distance settles at the network's own rounding
floor, ~1e-3 per evaluation (measured 9.6e-4 at Q4_K, T = 6000), the same size as this library's distance to the reference at F16
(4.5e-4).  NOISE = 3e-3 bounds those checks — an order of magnitude below the 2e-2 parity tolerance, two below what a wrong batch
index, a dropped CFG branch or a wrong dt produces (each test shows that with a negative control)."""
import numpy as np
import pytest

from tools import fixtures as fx
from parity_common import full_vae_dir, metrics, shift_schedule, turbo_dit_dir

pytestmark = pytest.mark.gpu

NOISE = 3e-3       # fp16 rounding floor of one 24-layer evaluation started from 1e-7-different fp32 values (see above)
EXACT = 0.0        # same kernels, same launch geometry, same rows
T240, T120, E = 6000, 3000, 512


@pytest.fixture(scope="module")
def br():
    from acestep_b200.bridge import B200Bridge
    b = B200Bridge()
    yield b
    b.close()


def batch_inputs(n, T, seed0=1234):
    inp = [fx.dit_inputs(fx.TURBO, T, E, seed=seed0 + i) for i in range(n)]
    return (np.stack([i["hidden_states"] for i in inp]), np.stack([i["context_latents"] for i in inp]),
            np.stack([i["encoder_hidden_states"] for i in inp]))


# ---------------------------------------------------------------- configs[2]: Q4_K, 240 s, full depth

@pytest.fixture(scope="module")
def q4k(br):
    d = turbo_dit_dir("Q4_K")
    br.load_dit(d)
    return d


def test_config2_batched_forward_equals_single_calls(br, q4k, monkeypatch):
    """3 songs at T = 6000 through ONE batched forward (M = 9000 rows per GEMM) against three ace_ggml_dit_forward calls (M = 3000)."""
    br.load_dit(q4k)
    h, c, e = batch_inputs(3, T240)
    ts = [1.0, 0.7, 0.3]
    vb = br.dit_forward_batched(h, c, e, None, None, ts, ts)
    assert np.isfinite(vb).all() and float(np.abs(vb).mean()) > 1e-3
    report = []
    singles = [br.dit_forward_tfirst(h[b], c[b], e[b], None, None, ts[b], ts[b]) for b in range(3)]
    for b in range(3):
        report.append(("batched vs single, default geometry", b, metrics(vb[b], singles[b])["rel_l2"], NOISE))
    report.append(("negative control: sample 1 of the batch against sample 0's single call", 1, metrics(vb[1], singles[0])["rel_l2"], None))
    # a sample's result does not depend on its neighbours: same sample 0, different company, same geometry -> bit-exact
    h2, c2, e2 = h.copy(), c.copy(), e.copy()
    h2[1:] *= -3.0
    e2[1:] = 0.0
    vb2 = br.dit_forward_batched(h2, c2, e2, None, None, ts, ts)
    report.append(("sample 0 with other neighbours (rel_l2)", 0, metrics(vb2[0], vb[0])["rel_l2"], EXACT))
    # K split pinned: batched and single run the same arithmetic per row -> bit-exact
    monkeypatch.setenv("ACE_B200_GEMM_SPLITK", "1")
    vp = br.dit_forward_batched(h, c, e, None, None, ts, ts)
    for b in range(3):
        v1 = br.dit_forward_tfirst(h[b], c[b], e[b], None, None, ts[b], ts[b])
        report.append(("batched vs single, K split pinned (rel_l2)", b, metrics(vp[b], v1)["rel_l2"], EXACT))
    print("\n".join(f"[property] {n} [{b}]: {v:.3e} (bound {bd})" for n, b, v, bd in report))
    assert report[3][2] > 0.5, report[3]
    assert all(v <= bd for _, _, v, bd in report if bd is not None), report


def test_config2_cfg_sampler_identities(br, q4k):
    """8-step shift-3 sampler, 2 songs at T = 6000: guidance 1 with a zero unconditional branch, and guidance 7 with the unconditional
    branch EQUAL to the conditional one, both reduce to the plain sampler (v_u + g (v_c - v_u) = v_c)."""
    br.load_dit(q4k)
    h, c, e = batch_inputs(2, T240)
    sch = br.shift_schedule(3.0, 8)
    assert np.allclose(sch, shift_schedule(3.0, 8), rtol=0, atol=1e-7)
    plain = br.generate_latents(h, c, e, None, schedule=sch, guidance=1.0)
    assert np.isfinite(plain).all()
    g1 = br.generate_latents(h, c, e, np.zeros_like(e), schedule=sch, guidance=1.0)
    m1 = metrics(g1, plain)["rel_l2"]           # v_u + (v_c - v_u) re-associates in fp32 -> rounding floor over 8 steps
    same = br.generate_latents(h, c, e, e.copy(), schedule=sch, guidance=7.0)
    m7 = metrics(same, plain)["rel_l2"]         # v_c - v_u = 0 exactly (same rows, same kernels); the 2B batch changes the K split
    # and the guided result differs from the plain one by far more than either bound (the combine is not a no-op)
    g7 = br.generate_latents(h, c, e, np.zeros_like(e), schedule=sch, guidance=7.0)
    mg = metrics(g7, plain)["rel_l2"]
    print(f"\n[property] CFG g=1 vs plain {m1:.3e}; identical branches g=7 vs plain {m7:.3e} (bound {NOISE}); negative control g=7 zero uncond {mg:.3e}")
    assert m1 <= NOISE and m7 <= NOISE and mg > 10 * NOISE, (m1, m7, mg)


def test_config2_device_sampler_equals_host_loop(br, q4k):
    """ace_b200_generate_latents (device-resident Euler loop) against the reference's host loop driven through ace_ggml_dit_forward
    (acestep_ggml.cpp:2056-2086: x -= v (t_i - t_{i+1}); last step x -= v t_i)."""
    br.load_dit(q4k)
    h, c, e = batch_inputs(1, T240, seed0=77)
    sch = br.shift_schedule(3.0, 8)
    dev = br.generate_latents(h, c, e, None, schedule=sch, guidance=1.0)[0]
    x = h[0].copy()
    for i, t in enumerate(sch):
        v = br.dit_forward_tfirst(x, c[0], e[0], None, None, float(t), float(t))
        dt = np.float32(t - sch[i + 1]) if i + 1 < len(sch) else np.float32(t)
        x = x - v * dt
    m = metrics(dev, x)["rel_l2"]
    print(f"\n[property] device sampler vs host loop: {m:.3e} (bound {NOISE})")
    assert m <= NOISE, m


def test_config2_padded_frames_do_not_leak(br, q4k):
    """T = 6000 with the last 1000 frames masked out: the velocities of the kept frames do not depend on what the padded frames hold."""
    br.load_dit(q4k)
    h, c, e = batch_inputs(1, T240, seed0=5)
    mask = np.ones((1, T240), np.int32)
    mask[:, 5000:] = 0
    va = br.dit_forward_batched(h, c, e, mask, None, [0.6], [0.6])
    h2, c2 = h.copy(), c.copy()
    h2[:, 5000:] = 11.0
    c2[:, 5000:, :64] = -7.0
    vb = br.dit_forward_batched(h2, c2, e, mask, None, [0.6], [0.6])
    assert np.isfinite(va).all()
    m = metrics(vb[:, :4990], va[:, :4990])   # frames 4990.. share a 2-frame patch / a conv tap with the padding
    assert m["rel_l2"] <= 1e-6, m


# ---------------------------------------------------------------- configs[4]: Q8_0, 120 s, 50-step closed-form schedule with CFG

def test_config4_50_step_cfg_sampler_equals_host_loop(br):
    """50 steps x (cond + uncond) at T = 3000 on the device against 100 single ace_ggml_dit_forward calls combined on the host."""
    br.load_dit(turbo_dit_dir("Q8_0"))
    h, c, e = batch_inputs(1, T120, seed0=21)
    sch = br.shift_schedule(3.0, 50)
    assert np.allclose(sch, shift_schedule(3.0, 50), rtol=0, atol=1e-7)
    un = np.zeros_like(e)
    dev = br.generate_latents(h, c, e, un, schedule=sch, guidance=7.0)[0]
    x = h[0].copy()
    for i, t in enumerate(sch):
        vc = br.dit_forward_tfirst(x, c[0], e[0], None, None, float(t), float(t))
        vu = br.dit_forward_tfirst(x, c[0], un[0], None, None, float(t), float(t))
        v = vu + np.float32(7.0) * (vc - vu)
        dt = np.float32(t - sch[i + 1]) if i + 1 < len(sch) else np.float32(t)
        x = x - v * dt
    m = metrics(dev, x)["rel_l2"]
    assert np.isfinite(dev).all()
    print(f"\n[property] 50-step CFG device sampler vs host loop: {m:.3e} (bound {(1 + 2 * 7.0) * NOISE:.1e})")
    # g = 7 amplifies each evaluation's rounding floor by (1 + 2 g) (the bound DESIGN.md states for guided sampling)
    assert m <= (1 + 2 * 7.0) * NOISE, m


# ---------------------------------------------------------------- configs[3]: VAE decode of one 600 s clip

@pytest.fixture(scope="module")
def clip600(br):
    br.load_vae(full_vae_dir())
    lat = fx.vae_latents(fx.VAE_FULL, 15000, seed=7)
    audio = br.vae_decode_tfirst(lat)
    assert audio.shape == (15000 * 1920, 2) and np.isfinite(audio).all()
    return lat, audio


@pytest.mark.parametrize("c0,c1", [(0, 1875), (1875, 3750), (7493, 7621), (13125, 15000)])
def test_config3_window_with_halo_equals_whole_clip(br, clip600, c0, c1):
    """core frames [c0, c1) decoded inside a window with a 16-frame halo (clip ends keep the decoder's own zero padding) equal the same
    samples of the un-split 600 s decode: the 8-GPU time split of configs[3] and the reference's chunk/overlap policy."""
    lat, audio = clip600
    br.load_vae(full_vae_dir())
    w0, w1 = max(0, c0 - 16), min(15000, c1 + 16)
    win = br.vae_decode_tfirst(lat[w0:w1])
    core = win[(c0 - w0) * 1920:(c1 - w0) * 1920]
    ref = audio[c0 * 1920:c1 * 1920]
    assert np.abs(core - ref).max() <= 1e-6 * max(1.0, float(np.abs(ref).max())), metrics(core, ref)


def test_config3_short_halo_is_detected(br, clip600):
    """the same comparison with a 4-frame halo (inside the 9.5-frame receptive field) must FAIL — the test above can see a wrong halo"""
    lat, audio = clip600
    br.load_vae(full_vae_dir())
    c0, c1 = 7493, 7621
    win = br.vae_decode_tfirst(lat[c0 - 4:c1 + 4])
    core = win[4 * 1920:(4 + c1 - c0) * 1920]
    assert np.abs(core - audio[c0 * 1920:c1 * 1920]).max() > 1e-4


def test_vae_240s_batch_members_are_independent_and_routes_agree(br, monkeypatch):
    """two 240 s clips decoded in one batched call equal their single decodes bit for bit, and the fused residual-unit kernel equals
    the two-launch route at full length (6000 frames: 45 000 / 90 000 row tiles on the two 128-channel levels)."""
    br.load_vae(full_vae_dir())
    lat = np.stack([fx.vae_latents(fx.VAE_FULL, 6000, seed=s) for s in (3, 4)])
    both = br.vae_decode_batched(lat)
    assert np.isfinite(both).all()
    for b in range(2):
        assert np.array_equal(both[b], br.vae_decode_tfirst(lat[b])), b
    monkeypatch.setenv("ACE_B200_VAE_FUSE", "0")
    plain = br.vae_decode_tfirst(lat[1])
    assert np.array_equal(plain, both[1]), metrics(plain, both[1])


# ---------------------------------------------------------------- run-to-run determinism of the quantised path at configs[2] sizes

def test_config2_evaluations_are_bit_reproducible(br, q4k, monkeypatch):
    """Identical calls give identical bits.  Before the decoders of the CTA-pair GEMM fenced their decoded sub-tile for the peer's tensor
    core (fence.proxy.async.shared::cluster; gemm_sm100.cu), a tensor core occasionally read a sub-tile that was not completely visible:
    13 distinct results in 60 single evaluations at M = 3000 and 8 in 12 calls of the 16-evaluation batch at M = 48 000, each off by
    1e-3 .. 2e-2 (profiles/r2_determinism_investigation.log).  30 single evaluations + 4 batched calls (64 evaluations) here."""
    import hashlib
    br.load_dit(q4k)
    h, c, e = batch_inputs(1, T240, seed0=1236)
    seen = {hashlib.md5(br.dit_forward_tfirst(h[0], c[0], e[0], None, None, 0.3, 0.3).tobytes()).hexdigest() for _ in range(30)}
    assert len(seen) == 1, f"{len(seen)} distinct outputs in 30 identical single evaluations"
    hb = np.concatenate([h] * 16); cb = np.concatenate([c] * 16); eb = np.concatenate([e] * 16)
    hb = hb * np.linspace(0.5, 1.5, 16, dtype=np.float32)[:, None, None]
    ts = [0.3 + 0.04 * i for i in range(16)]
    outs = [br.dit_forward_batched(hb, cb, eb, None, None, ts, ts) for _ in range(4)]
    assert np.isfinite(outs[0]).all()
    for o in outs[1:]:
        assert np.array_equal(o, outs[0]), metrics(o, outs[0])
    # the same with in-kernel decode at every size (ACE_B200_DECODE_PASS=0: the fenced CTA-pair decoders, the route of < 2048 rows)
    monkeypatch.setenv("ACE_B200_DECODE_PASS", "0")
    seen = {hashlib.md5(br.dit_forward_tfirst(h[0], c[0], e[0], None, None, 0.3, 0.3).tobytes()).hexdigest() for _ in range(30)}
    assert len(seen) == 1, f"in-kernel decode: {len(seen)} distinct outputs in 30 identical single evaluations"
    inner = [br.dit_forward_batched(hb, cb, eb, None, None, ts, ts) for _ in range(4)]
    for o in inner[1:]:
        assert np.array_equal(o, inner[0]), metrics(o, inner[0])
    m = metrics(inner[0], outs[0])   # the two routes feed the tensor cores the same fp16 operands
    assert m["rel_l2"] <= NOISE, m
