"""GPU: the allocation-free entry point (ace_b200_attention_noalloc, csrc/abi_noalloc.cu) against the existing one
(ace_b200_debug_attention) through the C ABI.  Collected LAST on purpose (a failure here hides nothing).

Same kernel, same fp16-rounded operands: the outputs are compared at the tolerance the attention tests use (rel 3e-3 against the
float64 restatement in test_kernels_gpu.py) — in practice they are bit-identical, which is asserted too.  Two calls in a row with
DIFFERENT shapes and a mask check that the shared static scratch array, which keeps the previous call's contents, leaks nothing
into the next result; a request beyond the 16 MiB array must fail with a clear message and leave the context usable.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def br():
    from acestep_b200.bridge import B200Bridge
    b = B200Bridge()
    yield b
    b.close()


def rel_err(a, ref):
    a, ref = np.asarray(a, np.float64), np.asarray(ref, np.float64)
    return float(np.linalg.norm(a - ref) / (np.linalg.norm(ref) + 1e-30))


def make_case(seed, B, ql, kl, nh, nkv, masked):
    g = np.random.default_rng(seed)
    Q = g.standard_normal((B, ql, nh * 128), dtype=np.float32)
    K = g.standard_normal((B, kl, nkv * 128), dtype=np.float32)
    V = g.standard_normal((B, kl, nkv * 128), dtype=np.float32)
    km = None
    if masked:
        km = (g.random((B, kl)) > 0.3).astype(np.int32)
        km[:, ::7] = 1
    return Q, K, V, km


def test_noalloc_attention_matches_the_allocating_entry_point_twice(br):
    # call 1: larger, unmasked, sliding window; call 2: smaller, masked, full attention (stale scratch beyond its extent)
    cases = [(make_case(1, 2, 300, 300, 4, 2, False), 4, 2, 128), (make_case(2, 1, 130, 77, 4, 2, True), 4, 2, -1)]
    for (Q, K, V, km), nh, nkv, window in cases:
        ref = br.debug_attention(Q, K, V, km, nh, nkv, window)
        out = br.attention_noalloc(Q, K, V, km, nh, nkv, window)
        assert out.shape == ref.shape and np.isfinite(out).all()
        assert rel_err(out, ref) < 3e-3, rel_err(out, ref)
        assert np.array_equal(out, ref)
    # and the first case again after the second: same answer as before
    (Q, K, V, km), nh, nkv, window = cases[0]
    assert np.array_equal(br.attention_noalloc(Q, K, V, km, nh, nkv, window), br.debug_attention(Q, K, V, km, nh, nkv, window))


def test_noalloc_attention_rejects_sizes_beyond_the_static_array(br):
    from acestep_b200.bridge import NOALLOC_SCRATCH_BYTES
    B, ql, kl, nh, nkv = 1, 2048, 2048, 8, 8   # nq = nk = 2 Mi elements -> 24 MiB of scratch > 16 MiB
    assert 4 * ql * nh * 128 * 3 > NOALLOC_SCRATCH_BYTES
    Q = np.zeros((B, ql, nh * 128), np.float32)
    K = np.zeros((B, kl, nkv * 128), np.float32)
    with pytest.raises(RuntimeError) as ei:
        br.attention_noalloc(Q, K, K, None, nh, nkv, -1)
    assert "scratch" in str(ei.value) and str(NOALLOC_SCRATCH_BYTES) in str(ei.value)
    # the context is still usable
    Qs, Ks, Vs, _ = make_case(3, 1, 64, 64, 2, 2, False)
    assert np.array_equal(br.attention_noalloc(Qs, Ks, Vs, None, 2, 2, -1), br.debug_attention(Qs, Ks, Vs, None, 2, 2, -1))
