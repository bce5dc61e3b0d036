"""GPU parity against the committed golden vectors (tests/golden/*.npz = outputs of the unmodified reference, made by
tests/golden/make_golden.py).  Unlike test_parity_gpu.py these need nothing from oracle/_ref at run time except, for the
K-quant GGUF fixtures, ggml_quantize_chunk (the same dependency the reference's own exporter has).
Tolerances as in test_parity_gpu.py."""
import os

import numpy as np
import pytest

from tools import fixtures as fx
from oracle import refbridge
from parity_common import make_dit_dir, make_vae_dir, metrics

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL = {"f32": (5e-3, 0.99999), "F16": (5e-3, 0.99999), "Q8_0": (2e-2, 0.9998), "Q6_K": (2e-2, 0.9998), "Q4_K": (2e-2, 0.9998)}


@pytest.fixture(scope="module")
def root(tmp_path_factory):
    return str(tmp_path_factory.mktemp("golden_fixtures"))


@pytest.fixture(scope="module")
def br():
    from acestep_b200.bridge import B200Bridge
    b = B200Bridge()
    yield b
    b.close()


@pytest.fixture(scope="module")
def dit_gold():
    return dict(np.load(os.path.join(GOLD, "dit_toy.npz")))


@pytest.fixture(scope="module")
def vae_gold():
    return dict(np.load(os.path.join(GOLD, "vae_toy.npz")))


@pytest.mark.parametrize("q", ["f32", "F16", "Q8_0", "Q4_K", "Q6_K"])
@pytest.mark.parametrize("route", ["in_kernel_decode", "single_cta_decode", "f16_twin", "decode_pass"])
def test_dit_forward_vs_golden(br, root, dit_gold, q, route, monkeypatch):
    if q in ("Q4_K", "Q6_K") and not refbridge.ref_available():
        pytest.fail("K-quant GGUF export needs ggml_quantize_chunk from oracle/_ref")
    if route != "in_kernel_decode" and q in ("f32", "F16"):
        pytest.skip("unquantised weights have one route")
    if route == "f16_twin":  # opt-in decoded twin, forced for every row count
        monkeypatch.setenv("ACE_B200_PREDECODE", "1")
        monkeypatch.setenv("ACE_B200_PREDECODE_MIN_M", "1")
    elif route == "single_cta_decode":
        monkeypatch.setenv("ACE_B200_GEMM_2CTA", "0")
    elif route == "decode_pass":  # per-launch decode into the stream's scratch + F16 kernels: the default from 2048 rows, forced here
        monkeypatch.setenv("ACE_B200_DECODE_PASS_MIN_M", "1")
    if route == "in_kernel_decode":
        monkeypatch.setenv("ACE_B200_DECODE_PASS", "0")
    d = make_dit_dir(root, fx.TOY, 0, None if q == "f32" else q)
    inp = fx.dit_inputs(fx.TOY, 33, 7, seed=1234)
    br.load_dit(d)
    out = br.dit_forward_tfirst(inp["hidden_states"], inp["context_latents"], inp["encoder_hidden_states"], None, None, 0.6, 0.6)
    m = metrics(out, dit_gold["v_" + q])
    rl, cs = TOL[q]
    assert m["rel_l2"] <= rl and m["cos"] >= cs, (q, route, m)


@pytest.mark.parametrize("k", [1, 2, 3, 4])
def test_dit_per_block_vs_golden(br, root, dit_gold, k, monkeypatch):
    d = make_dit_dir(root, fx.TOY, 0, None)
    inp = fx.dit_inputs(fx.TOY, 40, 9, seed=77)
    monkeypatch.setenv("ACE_GGML_DIT_MAX_LAYERS", str(k))
    br.load_dit(d)
    out = br.dit_forward_tfirst(inp["hidden_states"], inp["context_latents"], inp["encoder_hidden_states"], None, None, 0.3, 0.3)
    m = metrics(out, dit_gold[f"v_f32_layers{k}"])
    assert m["rel_l2"] <= 5e-3 and m["cos"] >= 0.99999, (k, m)


def test_dit_sampler_trace_vs_golden(br, root, dit_gold):
    d = make_dit_dir(root, fx.TOY, 0, None)
    inp = fx.dit_inputs(fx.TOY, 32, 8, seed=11)
    sched = br.shift_schedule(3.0, 8)
    br.load_dit(d)
    lat = br.generate_latents(inp["hidden_states"][None], inp["context_latents"][None], inp["encoder_hidden_states"][None], schedule=sched)
    assert metrics(lat[0], dit_gold["x_steps"][7])["rel_l2"] <= 2e-2


def test_dit_masks_vs_golden(br, root, dit_gold):
    d = make_dit_dir(root, fx.TOY, 0, None)
    inp = fx.dit_inputs(fx.TOY, 48, 12, seed=5)
    am = np.ones((48,), np.int32)
    am[40:] = 0
    em = np.ones((12,), np.int32)
    em[9:] = 0
    br.load_dit(d)
    out = br.dit_forward_tfirst(inp["hidden_states"], inp["context_latents"], inp["encoder_hidden_states"], am, em, 0.8, 0.5)
    assert metrics(out, dit_gold["v_masked"])["rel_l2"] <= 5e-3
    inp0 = fx.dit_inputs(fx.TOY, 48, 0, seed=6)
    out0 = br.dit_forward_tfirst(inp0["hidden_states"], inp0["context_latents"], None, None, None, 0.8, 0.8)
    assert metrics(out0, dit_gold["v_noenc"])["rel_l2"] <= 5e-3


@pytest.mark.parametrize("n", [1, 16, 45])
def test_vae_decode_vs_golden(br, root, vae_gold, n):
    d = make_vae_dir(root)
    br.load_vae(d)
    out = br.vae_decode_tfirst(fx.vae_latents(fx.VAE_TOY, n, seed=7))
    m = metrics(out, vae_gold[f"audio_{n}"])
    assert m["cos"] >= 0.9999 and m["snr_db"] >= 40.0, m


def test_vae_untiled_equals_reference_tiling_golden(br, root, vae_gold):
    d = make_vae_dir(root)
    br.load_vae(d)
    out = br.vae_decode_tfirst(fx.vae_latents(fx.VAE_TOY, 60, seed=9))
    m = metrics(out, vae_gold["audio_tiled_60"])
    assert m["cos"] >= 0.9999 and m["snr_db"] >= 40.0, m
