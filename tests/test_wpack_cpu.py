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
