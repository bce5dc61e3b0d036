"""GPU: the tap-GEMM (tcgen05) and attention kernels against numpy restatements, through the C ABI test hooks.

GEMM oracle: operands rounded to fp16 exactly as the kernel sees them (activations: cvt.rn; weights: the ggml
dequantisation of the same blocks, third_party/ggml/src/ggml-quants.c), products accumulated in float64.
Tolerance: fp32 accumulation order + (for quantised weights) the fp16 rounding of the decoded weight: 2e-3 relative
to the output RMS, stated per test.
"""
import numpy as np
import pytest

from blocks import random_rows

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def br():
    from acestep_b200.bridge import B200Bridge
    b = B200Bridge()
    yield b
    b.close()


def h16(x):
    return np.asarray(x, dtype=np.float32).astype(np.float16).astype(np.float64)


def tap_gemm_ref(A, W, taps, tap_stride, tap_offset, l_out):
    """A [B, L, C] fp64, W [N, taps, C] fp64 -> [B, l_out, N]"""
    B, L, C = A.shape
    N = W.shape[0]
    out = np.zeros((B, l_out, N))
    for t in range(taps):
        sh = t * tap_stride + tap_offset
        rows = np.arange(l_out) + sh
        ok = (rows >= 0) & (rows < L)
        At = np.zeros((B, l_out, C))
        At[:, ok] = A[:, rows[ok]]
        out += At @ W[:, t].T
    return out


def rel_err(a, b):
    return float(np.sqrt(np.mean((a - b) ** 2)) / (np.sqrt(np.mean(b ** 2)) + 1e-30))


def dequant_ref(rows, gt, N, K):
    from acestep_b200.bridge import pack_roundtrip
    _, ref = pack_roundtrip(rows, gt, N, K)
    return ref


@pytest.mark.parametrize("B,L,C,N", [(1, 128, 64, 128), (1, 200, 256, 384), (2, 130, 384, 200), (1, 1, 256, 512),
                                       (1, 700, 2048, 1024)])
@pytest.mark.parametrize("simt", [False, True])
def test_gemm_f16_linear(br, B, L, C, N, simt):
    g = np.random.default_rng(B * 7 + L + C + N)
    A = g.standard_normal((B, L, C), dtype=np.float32)
    W = (g.standard_normal((N, C), dtype=np.float32) * C ** -0.5).astype(np.float32)
    bias = g.standard_normal((N,), dtype=np.float32)
    out = br.debug_gemm(A, W, w_type=0, N=N, bias=bias, simt=simt)
    ref = tap_gemm_ref(h16(A), h16(W)[:, None, :], 1, 0, 0, L) + bias
    assert rel_err(out, ref) < 1e-4, rel_err(out, ref)


@pytest.mark.parametrize("gt", [8, 12, 14])
@pytest.mark.parametrize("L,K,N", [(128, 256, 128), (300, 512, 384), (64, 2048, 256)])
def test_gemm_quant_linear(br, gt, L, K, N):
    g = np.random.default_rng(gt * 31 + L + K + N)
    A = g.standard_normal((1, L, K), dtype=np.float32)
    rows = random_rows(gt, N, K, seed=gt + L)
    Wd = dequant_ref(rows, gt, N, K).astype(np.float64)
    ref = tap_gemm_ref(h16(A), Wd[:, None, :], 1, 0, 0, L)
    for simt in (True, False):
        out = br.debug_gemm(A, rows, w_type=gt, N=N, simt=simt)
        # decoded weights are rounded to fp16 in shared memory: <= 2^-11 relative per weight
        assert rel_err(out, ref) < 2e-3, (simt, rel_err(out, ref))


def test_gemm_q8_online(br):
    g = np.random.default_rng(5)
    A = g.standard_normal((1, 96, 384), dtype=np.float32)
    W = (g.standard_normal((256, 384), dtype=np.float32) * 0.05).astype(np.float32)
    out = br.debug_gemm(A, W, w_type=-8, N=256)
    ref = tap_gemm_ref(h16(A), W.astype(np.float64)[:, None, :], 1, 0, 0, 96)
    assert rel_err(out, ref) < 1e-2  # Q8_0 quantisation noise of the weights themselves


@pytest.mark.parametrize("dil", [1, 3, 9])
def test_gemm_conv7(br, dil):
    """k=7 dilated conv with zero padding = 7 taps, stride = dilation, offset = -3*dilation (acestep_vae_model.cpp:709)."""
    g = np.random.default_rng(dil)
    B, L, C, N = 2, 333, 64, 96
    A = g.standard_normal((B, L, C), dtype=np.float32)
    W = (g.standard_normal((N, 7, C), dtype=np.float32) * (7 * C) ** -0.5).astype(np.float32)
    out = br.debug_gemm(A, W, w_type=0, N=N, taps=7, tap_stride=dil, tap_offset=-3 * dil)
    ref = tap_gemm_ref(h16(A), h16(W), 7, dil, -3 * dil, L)
    assert rel_err(out, ref) < 1e-4


@pytest.mark.parametrize("dil,C", [(1, 128), (3, 128), (9, 128), (1, 64)])
def test_gemm_conv7_activation_slab(br, dil, C, monkeypatch):
    """persistent k7 convolution (>= 2 tiles per SM) through the activation-slab route — the tile's 128 + 6*dilation rows
    staged once, taps read at row offsets — against the per-k-block route: same MMAs in the same order, so bit-identical;
    and against the float64 restatement on the first / last rows (zero padding on both ends, ragged last tile) and a middle slice."""
    g = np.random.default_rng(100 + dil)
    B, L, N = 2, 148 * 128 + 77, 128
    A = g.standard_normal((B, L, C), dtype=np.float32)
    W = (g.standard_normal((N, 7, C), dtype=np.float32) * (7 * C) ** -0.5).astype(np.float32)
    bias = g.standard_normal((N,), dtype=np.float32)
    kw = dict(w_type=0, N=N, bias=bias, taps=7, tap_stride=dil, tap_offset=-3 * dil)
    monkeypatch.setenv("ACE_B200_GEMM_SLAB", "0")
    plain = br.debug_gemm(A, W, **kw)
    monkeypatch.setenv("ACE_B200_GEMM_SLAB", "1")
    slab = br.debug_gemm(A, W, **kw)
    assert np.array_equal(plain, slab)
    for lo, hi in ((0, 300), (L // 2, L // 2 + 300), (L - 300, L)):
        pad = 3 * dil
        a0, a1 = max(0, lo - pad), min(L, hi + pad)
        ref = tap_gemm_ref(h16(A[:, a0:a1]), h16(W), 7, dil, -3 * dil, a1 - a0)[:, lo - a0:hi - a0] + bias
        assert rel_err(slab[:, lo:hi], ref) < 1e-4, (lo, hi)


def test_gemm_conv_transpose_phases(br):
    """conv-transpose k=2s as 2 taps (rows i, i-1) x s phase columns (acestep_vae_model.cpp:696-707)."""
    g = np.random.default_rng(11)
    B, L, Cin, Cout, s = 1, 150, 64, 32, 4
    x = g.standard_normal((B, L, Cin), dtype=np.float32)
    w = (g.standard_normal((Cin, Cout, 2 * s), dtype=np.float32) * Cin ** -0.5).astype(np.float32)
    Wp = np.zeros((s * Cout, 2, Cin), dtype=np.float32)
    for r in range(s):
        for tap in range(2):
            Wp[r * Cout:(r + 1) * Cout, tap] = w[:, :, r + tap * s].T
    out = br.debug_gemm(x, Wp, w_type=0, N=s * Cout, taps=2, tap_stride=-1, tap_offset=0, l_out=L + 1)
    full = out.reshape(B, (L + 1) * s, Cout)
    xr, wr = h16(x), h16(w)
    ref = np.zeros((B, (L - 1) * s + 2 * s, Cout))
    for i in range(L):
        for k in range(2 * s):
            ref[:, i * s + k] += xr[:, i] @ wr[:, :, k]
    assert rel_err(full[:, : ref.shape[1]], ref) < 1e-4


def test_gemm_swiglu_and_residual_and_splitk(br):
    g = np.random.default_rng(2)
    L, K, I = 260, 512, 384
    A = g.standard_normal((1, L, K), dtype=np.float32)
    Wg = (g.standard_normal((I, K), dtype=np.float32) * K ** -0.5).astype(np.float32)
    Wu = (g.standard_normal((I, K), dtype=np.float32) * K ** -0.5).astype(np.float32)
    Wi = np.empty((2 * I, K), dtype=np.float32)
    Wi[0::2], Wi[1::2] = Wg, Wu
    out = br.debug_gemm(A, Wi, w_type=0, N=2 * I, swiglu=True)
    gg = h16(A)[0] @ h16(Wg).T
    uu = h16(A)[0] @ h16(Wu).T
    ref = gg / (1 + np.exp(-gg)) * uu
    assert rel_err(out[0], ref) < 2e-3  # fp16 output rounding
    # residual add, plain and split-K (atomics)
    W = (g.standard_normal((256, K), dtype=np.float32) * K ** -0.5).astype(np.float32)
    R = g.standard_normal((1, L, 256), dtype=np.float32)
    ref2 = h16(A)[0] @ h16(W).T + R[0]
    for sk in (False, True):
        o = br.debug_gemm(A, W, w_type=0, N=256, resid=R, split_k=sk)
        assert rel_err(o[0], ref2) < 1e-4, sk


def test_gemm_ragged_n_quant(br):
    """N = 640 = 5 tiles of 128 (odd count: the 128x256 variant must cope with a missing second sub-tile)."""
    g = np.random.default_rng(9)
    A = g.standard_normal((1, 512, 256), dtype=np.float32)
    rows = random_rows(12, 640, 256, seed=4)
    out = br.debug_gemm(A, rows, w_type=12, N=640)
    ref = tap_gemm_ref(h16(A), dequant_ref(rows, 12, 640, 256).astype(np.float64)[:, None, :], 1, 0, 0, 512)
    assert rel_err(out, ref) < 2e-3


def test_gemm_large_tiles(br):
    """enough tiles (>= 2 waves of 128x256) that the launcher picks BN=256 by itself, quantised + F16."""
    g = np.random.default_rng(10)
    L, K, N = 128 * 40, 256, 2048
    A = g.standard_normal((1, L, K), dtype=np.float32)
    rows = random_rows(8, N, K, seed=6)
    ref = tap_gemm_ref(h16(A), dequant_ref(rows, 8, N, K).astype(np.float64)[:, None, :], 1, 0, 0, L)
    out = br.debug_gemm(A, rows, w_type=8, N=N)
    assert rel_err(out, ref) < 2e-3
    W = (g.standard_normal((N, K), dtype=np.float32) * K ** -0.5).astype(np.float32)
    out = br.debug_gemm(A, W, w_type=0, N=N)
    assert rel_err(out, tap_gemm_ref(h16(A), h16(W)[:, None, :], 1, 0, 0, L)) < 1e-4


def attention_ref(Q, K, V, key_mask, n_heads, n_kv, window):
    B, ql, _ = Q.shape
    kl = K.shape[1]
    q = h16(Q).reshape(B, ql, n_heads, 128)
    k = h16(K).reshape(B, kl, n_kv, 128)
    v = h16(V).reshape(B, kl, n_kv, 128)
    rep = n_heads // n_kv
    out = np.zeros((B, ql, n_heads, 128))
    for b in range(B):
        for h in range(n_heads):
            s = q[b, :, h] @ k[b, :, h // rep].T / np.sqrt(128.0)
            allow = np.ones((ql, kl), dtype=bool)
            if window >= 0:
                d = np.arange(ql)[:, None] - np.arange(kl)[None, :]
                allow &= np.abs(d) <= window
            if key_mask is not None:
                allow &= (key_mask[b] != 0)[None, :]
            s = np.where(allow, s, -np.inf)
            s = s - s.max(axis=1, keepdims=True)
            p = np.exp(s)
            p /= p.sum(axis=1, keepdims=True)
            out[b, :, h] = p @ v[b, :, h // rep]
    return out.reshape(B, ql, n_heads * 128)


@pytest.mark.parametrize("impl", ["tcp", "tc4", "tc8", "mma"])  # P-in-TMEM (default) / smem-P 4 and 8 softmax warps / mma.sync
@pytest.mark.parametrize("ql,kl,window,masked", [(64, 64, -1, False), (200, 200, -1, False), (375, 375, 128, False),
                                                   (130, 77, -1, True), (300, 300, 16, True), (1, 5, -1, False),
                                                   (700, 700, -1, False), (1000, 1000, 128, False), (300, 512, -1, True)])
def test_attention(br, ql, kl, window, masked, impl, monkeypatch):
    monkeypatch.setenv("ACE_B200_ATTN", impl)
    g = np.random.default_rng(ql + kl)
    B, nh, nkv = 2, 4, 2
    Q = g.standard_normal((B, ql, nh * 128), dtype=np.float32)
    K = g.standard_normal((B, kl, nkv * 128), dtype=np.float32)
    V = g.standard_normal((B, kl, nkv * 128), dtype=np.float32)
    km = None
    if masked:
        km = (g.random((B, kl)) > 0.3).astype(np.int32)
        km[:, :: 7] = 1
        if window >= 0:  # keep every query's own key visible so no row is fully masked
            km[:] = 1
            km[:, 5:40] = 0
            window = max(window, 40)
    out = br.debug_attention(Q, K, V, km, nh, nkv, window)
    ref = attention_ref(Q, K, V, km, nh, nkv, window)
    # P and O are rounded to fp16 inside the kernel: ~1e-3 relative
    assert rel_err(out, ref) < 3e-3, rel_err(out, ref)


@pytest.mark.parametrize("gt", [8, 12, 14])
@pytest.mark.parametrize("shape", [(640, 512, 384), (129, 2048, 256), (1500, 1024, 1024), (3000, 512, 512)])
def test_gemm_quantised_routes_agree(br, gt, shape, monkeypatch):
    """quantised weight, three routes (gemm.h): decoded in shared memory by the single-CTA kernel, decoded in shared memory by the
    CTA-pair kernel (the default beyond one row tile; N = 384 has an odd tile count and falls back to the single CTA), and
    decoded once to the F16 twin (ACE_B200_PREDECODE=1) and run through the F16 kernels.  Same decoder arithmetic and the same K
    order per output element: the single-CTA route and the twin are bit-identical, the pair agrees to fp32 rounding."""
    L, K, N = shape
    g = np.random.default_rng(gt + L)
    A = g.standard_normal((1, L, K), dtype=np.float32)
    rows = random_rows(gt, N, K, seed=gt + 1)
    monkeypatch.setenv("ACE_B200_GEMM_2CTA", "0")
    single = br.debug_gemm(A, rows, w_type=gt, N=N)
    monkeypatch.delenv("ACE_B200_GEMM_2CTA")
    by_default = br.debug_gemm(A, rows, w_type=gt, N=N)   # >= 2048 rows: decode pass + F16 kernels, else in-kernel decode
    monkeypatch.setenv("ACE_B200_DECODE_PASS", "0")      # in-kernel decode at every size for the rest of the test
    default = br.debug_gemm(A, rows, w_type=gt, N=N)
    # publish mode of the pair kernel (ACE_B200_GEMM_QPUB=1): the decoded rows reach the operand tile through the bulk-copy engine
    # instead of generic stores + a cluster-scoped fence; same MMAs in the same order -> bit-identical to the default
    monkeypatch.setenv("ACE_B200_GEMM_QPUB", "1")
    published = br.debug_gemm(A, rows, w_type=gt, N=N)
    monkeypatch.delenv("ACE_B200_GEMM_QPUB")
    assert np.array_equal(published, default)
    monkeypatch.setenv("ACE_B200_PREDECODE", "1")
    monkeypatch.setenv("ACE_B200_PREDECODE_MIN_M", "1")
    monkeypatch.setenv("ACE_B200_GEMM_2CTA", "0")
    twin = br.debug_gemm(A, rows, w_type=gt, N=N)
    assert np.array_equal(single, twin)
    ref = tap_gemm_ref(h16(A), dequant_ref(rows, gt, N, K).astype(np.float64)[:, None, :], 1, 0, 0, L)
    assert rel_err(twin, ref) < 2e-3
    assert np.isfinite(default).all() and np.abs(default - single).max() <= 1e-5 * np.abs(ref).max(), np.abs(default - single).max()
    assert np.abs(by_default - single).max() <= 1e-5 * np.abs(ref).max(), np.abs(by_default - single).max()


@pytest.mark.parametrize("route", ["single", "pair"])
@pytest.mark.parametrize("wt", [0, 12])
def test_gemm_banded_rasterisation_is_a_permutation(br, route, wt, monkeypatch):
    """tile order in L2-sized bands of row tiles (tile_coords): with the band shrunk to a few row tiles — several bands, a short
    last one, two samples, K split — every output tile is still computed exactly once: bit-identical to the one-band order for
    plain stores, and to fp32 reduction order for the split-K accumulate epilogue."""
    g = np.random.default_rng(5)
    B, L, K, N = 2, 1100, 512, 768
    A = g.standard_normal((B, L, K), dtype=np.float32)
    W = (g.standard_normal((N, K), dtype=np.float32) * K ** -0.5) if wt == 0 else random_rows(wt, N, K, seed=3)
    R = g.standard_normal((B, L, N), dtype=np.float32)
    bias = g.standard_normal((N,), dtype=np.float32)
    monkeypatch.setenv("ACE_B200_GEMM_2CTA", "1" if route == "pair" else "0")
    for kw, exact in ((dict(bias=bias), True), (dict(resid=R), False), (dict(resid=R, split_k=True), False)):
        monkeypatch.setenv("ACE_B200_GEMM_BAND_MB", "0")
        one_band = br.debug_gemm(A, W, w_type=wt, N=N, **kw)
        outs = []
        for mb in ("0.3", "0.55", "0.8"):   # 128-row tiles of 128 KB / 256-row pairs of 256 KB: bands of 2..6 row tiles out of 10..18, short last bands
            monkeypatch.setenv("ACE_B200_GEMM_BAND_MB", mb)
            outs.append(br.debug_gemm(A, W, w_type=wt, N=N, **kw))
        for o in outs:
            if exact:
                assert np.array_equal(o, one_band)
            else:
                assert np.abs(o - one_band).max() <= 1e-5 * max(1.0, np.abs(one_band).max())


@pytest.mark.parametrize("gt", [8, 12, 14])
def test_gemm_quantised_pair_epilogues(br, gt, monkeypatch):
    """CTA-pair in-kernel decode through the fused epilogues the DiT uses: bias, SwiGLU over interleaved rows, split-K residual
    accumulate (TMA reduce-add); each against the single-CTA route"""
    g = np.random.default_rng(100 + gt)
    L, K, N = 700, 1024, 512
    A = g.standard_normal((1, L, K), dtype=np.float32)
    rows = random_rows(gt, N, K, seed=gt + 7)
    bias = g.standard_normal((N,), dtype=np.float32)
    R = g.standard_normal((1, L, N), dtype=np.float32)
    for kw in (dict(bias=bias), dict(swiglu=True), dict(resid=R), dict(resid=R, split_k=True)):
        monkeypatch.setenv("ACE_B200_GEMM_2CTA", "0")
        single = br.debug_gemm(A, rows, w_type=gt, N=N, **kw)
        monkeypatch.setenv("ACE_B200_GEMM_2CTA", "1")
        pair = br.debug_gemm(A, rows, w_type=gt, N=N, **kw)
        assert np.isfinite(pair).all()
        assert np.abs(pair - single).max() <= 2e-3 * max(1.0, np.abs(single).max()), (sorted(kw), np.abs(pair - single).max())


@pytest.mark.parametrize("shape", [(1, 700, 2048, 1024, 1), (2, 333, 128, 256, 7), (1, 128 * 20, 256, 2048, 1), (1, 100, 512, 512, 1)])
def test_gemm_cta_pair_matches_single_cta(br, shape, monkeypatch):
    """cta_group::2 route (256 x 256 per CTA pair) vs the single-CTA kernel on the same operands: same K order per
    output element, so results must agree to fp32 rounding; both against the float64 restatement."""
    B, L, C, N, taps = shape
    g = np.random.default_rng(L + N)
    A = g.standard_normal((B, L, C), dtype=np.float32)
    W = (g.standard_normal((N, taps, C), dtype=np.float32) * (taps * C) ** -0.5).astype(np.float32)
    bias = g.standard_normal((N,), dtype=np.float32)
    kw = dict(w_type=0, N=N, bias=bias, taps=taps, tap_stride=1 if taps > 1 else 0, tap_offset=-(taps // 2))
    monkeypatch.setenv("ACE_B200_GEMM_2CTA", "0")
    single = br.debug_gemm(A, W if taps > 1 else W[:, 0], **kw)
    monkeypatch.setenv("ACE_B200_GEMM_2CTA", "1")
    pair = br.debug_gemm(A, W if taps > 1 else W[:, 0], **kw)
    ref = tap_gemm_ref(h16(A), h16(W), taps, 1 if taps > 1 else 0, -(taps // 2), L) + bias
    assert rel_err(single, ref) < 1e-4
    assert rel_err(pair, ref) < 1e-4, rel_err(pair, ref)
    assert np.abs(pair - single).max() <= 1e-5 * np.abs(ref).max()


@pytest.mark.parametrize("wt", [0, 12])
def test_gemm_cta_pair_stress_is_deterministic(br, wt, monkeypatch):
    """the CTA-pair kernels publish "CTA 1's stage landed" to the leader with a RELAXED remote mbarrier arrive (ptx.cuh: a release at
    cluster scope cost ~1 us per k-block) — for the QUANT variant the stage was written by generic stores + fence.proxy.async.  A lost
    ordering would show as run-to-run differences: 25 repetitions x 3 shapes must be bit-identical to the first run of the same
    route, and agree with the single-CTA route."""
    g = np.random.default_rng(77)
    for (L, K, N) in ((2100, 2048, 1024), (900, 6144, 512), (5000, 512, 2048)):
        A = g.standard_normal((1, L, K), dtype=np.float32)
        W = (g.standard_normal((N, K), dtype=np.float32) * K ** -0.5) if wt == 0 else random_rows(wt, N, K, seed=L)
        monkeypatch.setenv("ACE_B200_GEMM_2CTA", "0")
        single = br.debug_gemm(A, W, w_type=wt, N=N)
        monkeypatch.setenv("ACE_B200_GEMM_2CTA", "1")
        first = br.debug_gemm(A, W, w_type=wt, N=N)
        assert np.abs(first - single).max() <= 1e-5 * np.abs(single).max()
        for it in range(24):
            again = br.debug_gemm(A, W, w_type=wt, N=N)
            assert np.array_equal(again, first), (L, K, N, it)


def test_gemm_cta_pair_splitk_residual(br, monkeypatch):
    g = np.random.default_rng(21)
    L, K, N = 260, 1024, 512
    A = g.standard_normal((1, L, K), dtype=np.float32)
    W = (g.standard_normal((N, K), dtype=np.float32) * K ** -0.5).astype(np.float32)
    R = g.standard_normal((1, L, N), dtype=np.float32)
    ref = h16(A)[0] @ h16(W).T + R[0]
    monkeypatch.setenv("ACE_B200_GEMM_2CTA", "1")
    for sk in (False, True):
        o = br.debug_gemm(A, W, w_type=0, N=N, resid=R, split_k=sk)
        assert rel_err(o[0], ref) < 1e-4, sk


@pytest.mark.parametrize("impl", ["tcp", "tc4", "tc8"])
def test_attention_lazy_rescale(br, impl, monkeypatch):
    """scores that grow by far more than 2^8 along the key axis force the tcgen05 kernel through its O / l rescale path
    (stale reference max) several times; the result must not depend on it"""
    monkeypatch.setenv("ACE_B200_ATTN", impl)
    g = np.random.default_rng(3)
    B, nh, nkv, ql, kl = 1, 2, 1, 256, 640
    Q = g.standard_normal((B, ql, nh * 128), dtype=np.float32)
    K = g.standard_normal((B, kl, nkv * 128), dtype=np.float32)
    K *= np.linspace(0.2, 6.0, kl, dtype=np.float32)[None, :, None]  # later keys score much higher in magnitude
    V = g.standard_normal((B, kl, nkv * 128), dtype=np.float32)
    out = br.debug_attention(Q, K, V, None, nh, nkv, -1)
    ref = attention_ref(Q, K, V, None, nh, nkv, -1)
    assert rel_err(out, ref) < 3e-3, rel_err(out, ref)
