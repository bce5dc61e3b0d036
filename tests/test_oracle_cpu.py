"""CPU: pins the NumPy restatement (oracle/dit_numpy.py, oracle/vae_numpy.py) against golden vectors produced by the
UNMODIFIED reference built from source (tests/golden/make_golden.py -> tests/golden/*.npz).

The reference tree asserts no known answers for this path ("In-tree: none asserted"), so the goldens are
outputs of the reference itself run in the build container.  Tolerances:
  * F32 weights: the reference computes f32 dot products, the restatement float64 -> rel-L2 <= 2e-5
  * VAE: both sides round conv inputs/weights to fp16 at the same points; f32 vs f64 accumulation flips a few fp16 roundings in
    the 30-conv stack -> rel-L2 <= 2e-3 (65 dB SNR measured; the GPU tolerance is 40 dB)
  * quantised goldens vs the F32 truth: the reference's own quantisation noise (weights + int8 activations), bounded by the
    SURVEY 8c tolerances that the GPU parity tests use
"""
import os

import numpy as np
import pytest

from tools import fixtures as fx
from oracle import dit_numpy as dn
from oracle import vae_numpy as vn
from parity_common import metrics

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def dit_gold():
    return dict(np.load(os.path.join(GOLD, "dit_toy.npz")))


@pytest.fixture(scope="module")
def vae_gold():
    return dict(np.load(os.path.join(GOLD, "vae_toy.npz")))


@pytest.fixture(scope="module")
def W():
    return fx.dit_tensors(fx.TOY, 0), fx.TOY.config_json()


@pytest.fixture(scope="module")
def VW():
    return fx.vae_tensors(fx.VAE_TOY, 0), fx.VAE_TOY.config_json()


def test_dit_full_forward_f32(W, dit_gold):
    T, cfg = W
    inp = fx.dit_inputs(fx.TOY, 33, 7, seed=1234)
    v = dn.forward_dit(T, cfg, inp["hidden_states"], inp["context_latents"], inp["encoder_hidden_states"], None, None, 0.6, 0.6)
    m = metrics(v, dit_gold["v_f32"])
    assert m["rel_l2"] <= 2e-5, m


@pytest.mark.parametrize("k", [1, 2, 3, 4])
def test_dit_per_block_f32(W, dit_gold, k):
    T, cfg = W
    inp = fx.dit_inputs(fx.TOY, 40, 9, seed=77)
    v = dn.forward_dit(T, cfg, inp["hidden_states"], inp["context_latents"], inp["encoder_hidden_states"], None, None, 0.3, 0.3, max_layers=k)
    m = metrics(v, dit_gold[f"v_f32_layers{k}"])
    assert m["rel_l2"] <= 2e-5, (k, m)


def test_dit_masks_tr_and_no_encoder(W, dit_gold):
    T, cfg = W
    inp = fx.dit_inputs(fx.TOY, 48, 12, seed=5)
    am = np.ones((48,), np.int32)
    am[40:] = 0
    em = np.ones((12,), np.int32)
    em[9:] = 0
    v = dn.forward_dit(T, cfg, inp["hidden_states"], inp["context_latents"], inp["encoder_hidden_states"], am, em, 0.8, 0.5)
    assert metrics(v, dit_gold["v_masked"])["rel_l2"] <= 2e-5
    inp0 = fx.dit_inputs(fx.TOY, 48, 0, seed=6)
    v0 = dn.forward_dit(T, cfg, inp0["hidden_states"], inp0["context_latents"], None, None, None, 0.8, 0.8)
    assert metrics(v0, dit_gold["v_noenc"])["rel_l2"] <= 2e-5


def test_dit_euler_trace(W, dit_gold):
    T, cfg = W
    inp = fx.dit_inputs(fx.TOY, 32, 8, seed=11)
    sched = dn.shift_schedule(3.0, 8)
    _, steps = dn.euler_sample(T, cfg, inp["hidden_states"], inp["context_latents"], inp["encoder_hidden_states"], sched)
    for i in range(8):
        assert metrics(steps[i], dit_gold["x_steps"][i])["rel_l2"] <= 5e-5, i


def test_shift_schedule_reproduces_reference_tables():
    """acestep_ggml.cpp:1485-1487"""
    s3 = [1.0, 0.9545454545, 0.9, 0.8333333333, 0.75, 0.6428571429, 0.5, 0.3]
    s2 = [1.0, 0.9333333333, 0.8571428571, 0.7692307692, 0.6666666667, 0.5454545455, 0.4, 0.2222222222]
    assert np.allclose(dn.shift_schedule(3.0, 8), s3, atol=1e-7)
    assert np.allclose(dn.shift_schedule(2.0, 8), s2, atol=1e-7)
    assert np.allclose(dn.shift_schedule(1.0, 8), 1.0 - np.arange(8) / 8, atol=1e-7)


@pytest.mark.parametrize("q,rl,cs", [("F16", 5e-3, 0.99999), ("Q8_0", 3e-2, 0.9995), ("Q6_K", 8e-2, 0.997), ("Q4_K", 2.5e-1, 0.97)])
def test_reference_quant_noise_vs_f32_truth(W, dit_gold, q, rl, cs):
    """documents how far the reference's own quantised paths sit from the UNQUANTISED F32 truth on this fixture (weight
    quantisation + int8 activation re-quantisation): measured rel-L2 9.6e-4 / 1.9e-2 / 4.9e-2 / 1.9e-1 for F16/Q8_0/Q6_K/Q4_K"""
    T, cfg = W
    inp = fx.dit_inputs(fx.TOY, 33, 7, seed=1234)
    v = dn.forward_dit(T, cfg, inp["hidden_states"], inp["context_latents"], inp["encoder_hidden_states"], None, None, 0.6, 0.6)
    m = metrics(dit_gold[f"v_{q}"], v)
    assert m["rel_l2"] <= rl and m["cos"] >= cs, (q, m)


@pytest.mark.parametrize("n", [1, 16, 45])
def test_vae_decode(VW, vae_gold, n):
    T, cfg = VW
    lat = fx.vae_latents(fx.VAE_TOY, n, seed=7)
    a = vn.decode(T, cfg, lat)
    g = vae_gold[f"audio_{n}"]
    assert a.shape == g.shape == (n * fx.VAE_TOY.hop, 2)
    m = metrics(a, g)
    assert m["rel_l2"] <= 2e-3, m


def test_vae_tiled_policy(VW, vae_gold):
    """chunk 36 / overlap 12 (halo >= receptive field): tiled == golden tiled, and == un-tiled up to fp16 boundary rounding"""
    T, cfg = VW
    lat = fx.vae_latents(fx.VAE_TOY, 60, seed=9)
    t = vn.decode_tiled(T, cfg, lat, 36, 12)
    assert metrics(t, vae_gold["audio_tiled_60"])["rel_l2"] <= 2e-3
    u = vn.decode(T, cfg, lat)
    assert metrics(u, vae_gold["audio_tiled_60"])["rel_l2"] <= 4e-3


@pytest.mark.parametrize("q,rl", [("F16", 2e-3), ("Q8_0", 2e-2), ("Q6_K", 2e-2), ("Q4_K", 2e-2)])
def test_reference_activation_noise_vs_dequantised_truth(dit_gold, q, rl, tmp_path):
    """same QUANTISED weights on both sides (GGUF twin dequantised with gguf.quants): what remains is the reference's
    activation re-quantisation (int8 per 32 / 256 block; fp16 for F16 weights; ggml-cpu.c:1286-1324) — the quantity the GPU tolerance is about"""
    from parity_common import dequantised_gguf_tensors
    d = str(tmp_path / "fx")
    lib = None
    if q in ("Q6_K", "Q4_K"):
        from oracle import refbridge
        if not refbridge.ref_available():
            pytest.skip("K-quant export needs ggml_quantize_chunk from oracle/_ref")
        lib = refbridge.ref_lib()
    fx.make_dit_fixture(d, fx.TOY, 0, safetensors=False, gguf_quant=q, ggml_lib=lib)
    T = dequantised_gguf_tensors(os.path.join(d, "model.gguf"))
    inp = fx.dit_inputs(fx.TOY, 33, 7, seed=1234)
    v = dn.forward_dit(T, fx.TOY.config_json(), inp["hidden_states"], inp["context_latents"], inp["encoder_hidden_states"], None, None, 0.6, 0.6)
    m = metrics(dit_gold[f"v_{q}"], v)
    assert m["rel_l2"] <= rl, (q, m)
