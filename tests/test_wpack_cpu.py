"""CPU: the tile packer + the host emulation of the kernel's decoder reproduce the ggml dequantisation
(reference formulas: ggml-quants.c:401-416 q8_0, 1352-1375 q4_K, 1762-1792 q6_K)."""
import numpy as np
import pytest

from acestep_b200 import bridge
from blocks import random_rows


@pytest.mark.parametrize("gt,K", [(8, 64), (8, 384), (8, 2048), (12, 256), (12, 1024), (14, 256), (14, 768)])
@pytest.mark.parametrize("N", [1, 100, 128, 300])
def test_pack_roundtrip_quant(gt, K, N):
    rows = random_rows(gt, N, K, seed=gt * 1000 + N + K)
    un, ref = bridge.pack_roundtrip(rows, gt, N, K)
    assert np.isfinite(ref).all()
    # folded scales are rounded to fp16 (2^-11 relative) and the product is formed in fp32 on the host emulation
    tol = 2.0 ** -10 * np.abs(ref).max() + 1e-12
    assert np.abs(un - ref).max() <= tol


def test_pack_roundtrip_f16():
    g = np.random.default_rng(3)
    W = g.standard_normal((200, 192), dtype=np.float32)
    un, ref = bridge.pack_roundtrip(W, 0, 200, 192)
    assert np.array_equal(un, ref)
    assert np.array_equal(ref, W.astype(np.float16).astype(np.float32))


@pytest.mark.parametrize("gt,blk_bytes", [(8, 34 * 8), (12, 144), (14, 210)])
def test_online_quantisers_are_bit_identical_to_ggml(gt, blk_bytes):
    """the row quantisers ACE_GGML_*_WEIGHT_QTYPE runs at load (wpack.cpp: quantize_row_q8_0 / q4_k / q6_k) against ggml_quantize_chunk
    of the reference library (what the reference runs, acestep_dit_model.cpp:156-192): every byte equal, on ordinary rows and on the
    corner cases of the fits (all-zero groups, constant groups, one outlier, all-positive / all-negative rows, tiny magnitudes)"""
    import ctypes
    from oracle import refbridge
    if not refbridge.ref_available():
        pytest.skip("needs ggml_quantize_chunk from oracle/_ref")
    lib = bridge.load_library()
    ref = ctypes.CDLL(refbridge.ref_lib())
    ref.ggml_quantize_chunk.argtypes = [ctypes.c_int, ctypes.POINTER(ctypes.c_float), ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64,
                                        ctypes.c_int64, ctypes.POINTER(ctypes.c_float)]
    ref.ggml_quantize_chunk.restype = ctypes.c_size_t
    g = np.random.default_rng(gt)
    N, K = 96, 1024
    W = (g.standard_normal((N, K)) * 0.03).astype(np.float32)
    W[0] = 0.0
    W[1] = 0.25
    W[2, :256] = 0.0
    W[3] = np.abs(W[3])
    W[4] = -np.abs(W[4])
    W[5, 17] = 40.0
    W[6] *= 1e-20
    W[7] = (g.standard_normal(K) * 3e4).astype(np.float32)
    W[8, ::2] = 0.0
    W[9] = np.round(W[9] * 100) / 100
    rb = K // 256 * blk_bytes
    mine = np.zeros((N, rb), np.uint8)
    theirs = np.zeros((N, rb), np.uint8)
    F = ctypes.POINTER(ctypes.c_float)
    assert lib.ace_b200_debug_quantize_rows(W.ctypes.data_as(F), gt, N, K, mine.ctypes.data_as(ctypes.c_void_p), mine.nbytes) == 0
    assert ref.ggml_quantize_chunk(gt, W.ctypes.data_as(F), theirs.ctypes.data_as(ctypes.c_void_p), 0, N, K, None) == theirs.nbytes
    bad = np.nonzero((mine != theirs).any(axis=1))[0]
    assert bad.size == 0, (gt, bad[:10])
